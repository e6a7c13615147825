mkdir -p gpurun_out
timeout 1700 python -m pytest tests/test_gpu_pair_tiles.py -m gpu -x -q 2>&1 | tail -15
