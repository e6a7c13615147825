mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_strided_bwd.py -m gpu -q --maxfail=6 > gpurun_out/r02_tests_a.log 2>&1; echo "rc=$?"
grep -n "^E  .*\|passed\|failed\|^FAILED" gpurun_out/r02_tests_a.log | cut -c1-260 | head -30
timeout 600 python bench.py --model alexnet --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/tmp_bench.json 2> gpurun_out/tmp_bench.err
python -c "
import json
j=json.load(open('gpurun_out/tmp_bench.json')); print('alexnet', j['value'], j['ms_per_step'], j['e2e']['value'], j['roofline']['frac'])
"; tail -2 gpurun_out/tmp_bench.err
