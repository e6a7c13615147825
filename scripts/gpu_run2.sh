mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_replay.py -m gpu -q -s --maxfail=20 > gpurun_out/r02_tests_replay.log 2>&1; echo "replay rc=$?" >> gpurun_out/r02_tests_replay.log
grep -n "AssertionError\|passed\|failed" gpurun_out/r02_tests_replay.log | head -30
grep -n "^tiny\|^lenet\|^vgg\|^alexnet\|^resnet\|^simple" gpurun_out/r02_tests_replay.log | cut -c1-1500
timeout 300 python scripts/e2e_debug.py vgg16 2>&1 | tail -8
