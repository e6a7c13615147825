mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --maxfail=10 -x > gpurun_out/r02_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_tests.log
grep -n "AssertionError\|passed\|failed\|Error" gpurun_out/r02_tests.log | head -30
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02d_bench_resnet50.json 2> gpurun_out/r02d_bench_resnet50.err
python -c "
import json,sys
j=json.load(open('gpurun_out/r02d_bench_resnet50.json')); print('resnet50', j['value'], j['ms_per_step'], j['e2e']['value'], j['roofline']['frac'], j['roofline']['achieved'], j['gpu_launches'])
"
tail -3 gpurun_out/r02d_bench_resnet50.err
timeout 600 python scripts/timeline.py resnet50 gpurun_out/r02_timeline_resnet50.json 2>&1 | tail -1
python scripts/timeline_report.py gpurun_out/r02_timeline_resnet50.json > gpurun_out/r02_timeline_report.txt; head -8 gpurun_out/r02_timeline_report.txt; tail -16 gpurun_out/r02_timeline_report.txt
