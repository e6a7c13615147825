mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_parallel.py -m gpu -q -x > gpurun_out/r02_tests_dp.log 2>&1; echo "rc=$?" >> gpurun_out/r02_tests_dp.log
tail -5 gpurun_out/r02_tests_dp.log; for r in 0 1; do echo "--- rank $r"; tail -40 gpurun_out/dp_rank$r.log; done
