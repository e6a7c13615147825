mkdir -p gpurun_out
timeout 400 python scripts/pair_bench.py 2>&1 | tail -16
