mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --maxfail=10 > gpurun_out/r02_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_tests.log
grep -n "AssertionError\|passed\|failed\|Error" gpurun_out/r02_tests.log | head -30
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --dump-profile gpurun_out/r02b_prof_resnet50.json > gpurun_out/r02b_bench_resnet50.json 2> gpurun_out/r02b_bench_resnet50.err
timeout 600 python bench.py --model vgg16 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02b_bench_vgg16.json 2> gpurun_out/r02b_bench_vgg16.err
for m in resnet50 vgg16; do python -c "
import json,sys
try:
    j=json.load(open('gpurun_out/r02b_bench_$m.json')); print('$m', j['value'], j['ms_per_step'], j['e2e']['value'], j['roofline']['frac'], j['roofline']['achieved'], j['hbm_kernels'])
except Exception as e: print('$m', 'FAILED', e)
"; done
tail -3 gpurun_out/r02b_bench_resnet50.err
