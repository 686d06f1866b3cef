mkdir -p gpurun_out
A="16 12 128 4096 1 0 bf16"; B="32 12 128 2048 1 0 bf16"; C="8 12 64 4096 1 0 fp16"; D="8 12 128 8192 1 0 bf16"
timeout 900 python tools/time_many.py \
 "$A FA_BH_GROUP=16" "$A FA_BH_GROUP=24" "$A FA_BH_GROUP=20" "$A FA_BH_GROUP=16" "$A FA_BH_GROUP=24" "$A FA_BH_GROUP=20" "$A FA_BH_GROUP=16" "$A FA_BH_GROUP=24" \
 "$B FA_BH_GROUP=24" "$B FA_BH_GROUP=32" "$B FA_BH_GROUP=48" "$B FA_BH_GROUP=64" "$B FA_BH_GROUP=24" "$B FA_BH_GROUP=32" "$B FA_BH_GROUP=48" "$B FA_BH_GROUP=64" \
 "$C FA_BH_GROUP=12" "$C FA_BH_GROUP=16" "$C FA_BH_GROUP=24" "$C FA_BH_GROUP=32" "$C FA_BH_GROUP=12" "$C FA_BH_GROUP=16" "$C FA_BH_GROUP=24" "$C FA_BH_GROUP=32" \
 "$D FA_BH_GROUP=6" "$D FA_BH_GROUP=8" "$D FA_BH_GROUP=12" "$D FA_BH_GROUP=6" "$D FA_BH_GROUP=8" "$D FA_BH_GROUP=12" \
 > gpurun_out/c4_time_many.log 2>&1
tail -32 gpurun_out/c4_time_many.log
