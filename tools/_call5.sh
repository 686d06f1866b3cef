mkdir -p gpurun_out
( time timeout 2400 python -m pytest tests -m gpu -x -q ) > gpurun_out/c5_pytest_gpu.log 2>&1
tail -5 gpurun_out/c5_pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/c5_bench.json 2> gpurun_out/c5_bench.err
cat gpurun_out/c5_bench.json
