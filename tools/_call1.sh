set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/c2_smi.txt 2>&1
timeout 600 python -m pytest tests/test_features_gpu.py -x -q -k "rotary or mha" > gpurun_out/c2_rotary.log 2>&1
timeout 900 python tools/time_many.py \
 "16 12 128 4096 1 0 bf16 FA_BH_GROUP=24" "16 12 128 4096 1 0 bf16 FA_BH_GROUP=12" "16 12 128 4096 1 0 bf16 FA_BH_GROUP=8" \
 "8 12 128 8192 1 0 bf16 FA_BH_GROUP=24" "8 12 128 8192 1 0 bf16 FA_BH_GROUP=12" "8 12 128 8192 1 0 bf16 FA_BH_GROUP=6" "8 12 128 8192 1 0 bf16 FA_BH_GROUP=3" \
 "4 12 128 16384 1 0 bf16 FA_BH_GROUP=24" "4 12 128 16384 1 0 bf16 FA_BH_GROUP=12" "4 12 128 16384 1 0 bf16 FA_BH_GROUP=6" "4 12 128 16384 1 0 bf16 FA_BH_GROUP=3" "4 12 128 16384 1 0 bf16 FA_BH_GROUP=1" \
 "16 12 128 4096 0 0 bf16" "16 12 128 4096 1 0.1 bf16" \
 "8 12 64 4096 0 0 fp16" "8 12 64 4096 0 0.1 fp16" "8 12 64 4096 1 0 fp16" "8 12 64 4096 1 0.1 fp16" \
 "8 12 64 2048 1 0 fp16" "8 12 64 2048 1 0.1 fp16" "8 12 64 1024 1 0 fp16" "8 12 64 1024 1 0.1 fp16" "8 12 64 512 0 0 fp16" \
 "32 12 64 4096 1 0.1 fp16" "32 12 64 1024 1 0.1 fp16" \
 > gpurun_out/c2_time_many.log 2>&1
FA_B200_LIB=$PWD/hazyresearch_flash-attention_b200/csrc/libfa_b200_trace.so timeout 300 python tools/trace_bwd.py 16 12 128 4096 1 > gpurun_out/c2_trace_causal.log 2>&1
FA_B200_LIB=$PWD/hazyresearch_flash-attention_b200/csrc/libfa_b200_trace.so timeout 300 python tools/trace_bwd.py 8 12 64 4096 0 > gpurun_out/c2_trace_d64.log 2>&1
timeout 300 true > gpurun_out/c2_host.log 2>&1
tail -30 gpurun_out/c2_time_many.log
