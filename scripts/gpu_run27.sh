mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_helpers.py tests/test_gpu_bn_maxpool.py tests/test_zz_no_alloc.py tests/test_gpu_bwd_bnstats.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>&1 | tail -1 | tee gpurun_out/r27_bench.json | cut -c1-200
