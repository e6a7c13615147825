mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --maxfail=15 > gpurun_out/r02_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_tests.log
grep -n "^E  .*Error\|passed\|failed\|^FAILED" gpurun_out/r02_tests.log | cut -c1-400 | head -40
