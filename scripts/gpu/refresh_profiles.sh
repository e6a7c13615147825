mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q --maxfail=10 > gpurun_out/r02_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_tests.log
tail -4 gpurun_out/r02_tests.log
# launch list of one steady-state step (plain run first, then the same command under ncu)
python scripts/one_step.py 1 1 > gpurun_out/one_step_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r02_launches_all.csv python scripts/one_step.py 1 1 > gpurun_out/ncu1.log 2>&1
echo "launch list rc=$?"
python scripts/agg_launches.py gpurun_out/r02_launches_all.csv --last-step k_pack_stem_filter --out gpurun_out/r02_launches_step.csv > gpurun_out/r02_launches_step.summary.txt; head -3 gpurun_out/r02_launches_step.summary.txt
python scripts/conv_traffic.py gpurun_out/r02_launches_step.csv gpurun_out/r02_conv_traffic.json
rm -f gpurun_out/r02_launches_all.csv
cp gpurun_out/r02_conv_traffic.json profiles/r02_conv_traffic.json
for m in resnet50 vgg16 alexnet lenet; do
  timeout 900 python bench.py --model $m --steps 20 --warmup 5 --dump-profile gpurun_out/r02_prof_$m.json > gpurun_out/r02_bench_$m.json 2> gpurun_out/r02_bench_$m.err
  python -c "
import json,sys
try:
    j=json.load(open('gpurun_out/r02_bench_$m.json')); print('$m', round(j['value']), round(j['ms_per_step'],3), round(j['e2e']['value']), round(j['roofline']['frac'],3), round(j['roofline']['achieved'],1), j['gpu_launches'], j.get('cpu_baseline',{}).get('value'), j['clocks'])
except Exception as e: print('$m', 'FAILED', e)
"
done
for m in resnet50 vgg16 alexnet lenet; do
  timeout 600 python bench.py --impl reference --model $m --steps 3 --warmup 1 > gpurun_out/r02_bench_reference_$m.json 2> gpurun_out/r02_bench_reference_$m.err
  python -c "
import json
j=json.load(open('gpurun_out/r02_bench_reference_$m.json')); print('ref $m', j['value'], j['ms_per_step'], j['cpu_baseline']['cores'])
"
done
timeout 600 python scripts/timeline.py resnet50 gpurun_out/r02_timeline_resnet50.json 2>&1 | tail -1
python scripts/timeline_report.py gpurun_out/r02_timeline_resnet50.json > gpurun_out/r02_timeline_report.txt; head -5 gpurun_out/r02_timeline_report.txt
