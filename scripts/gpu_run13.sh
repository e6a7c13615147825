mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --maxfail=10 > gpurun_out/r02_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_tests.log
grep -n "^E  .*Error\|passed\|failed\|^FAILED" gpurun_out/r02_tests.log | cut -c1-300 | head -30
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --dump-profile gpurun_out/r02f_prof_resnet50.json > gpurun_out/r02f_bench_resnet50.json 2> gpurun_out/r02f_bench_resnet50.err
python -c "
import json,sys
j=json.load(open('gpurun_out/r02f_bench_resnet50.json')); print('resnet50', j['value'], j['ms_per_step'], j['e2e']['value'], j['roofline']['frac'], j['roofline']['achieved'], j['gpu_launches'], {k:(round(v['achieved_gbs']),round(v['share_of_step'],3)) for k,v in j['hbm_kernels'].items()})
"
tail -3 gpurun_out/r02f_bench_resnet50.err
