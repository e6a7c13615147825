mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_bn_maxpool.py -m gpu -x -q 2>&1 | tail -15
timeout 900 python -m pytest tests/test_gpu_replay.py tests/test_gpu_network.py -m gpu -x -q 2>&1 | tail -15
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>&1 | tail -2 | tee gpurun_out/r23_bench.json
