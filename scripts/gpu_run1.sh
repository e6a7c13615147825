mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r02_gpuinfo.txt 2>&1
nproc >> gpurun_out/r02_gpuinfo.txt; free -g | head -2 >> gpurun_out/r02_gpuinfo.txt
timeout 1500 python -m pytest tests -m gpu -q --maxfail=12 -x --deselect tests/test_gpu_replay.py > gpurun_out/r02_tests_old.log 2>&1; echo "old tests rc=$?" >> gpurun_out/r02_tests_old.log
timeout 1500 python -m pytest tests/test_gpu_replay.py -m gpu -q -s --maxfail=20 > gpurun_out/r02_tests_replay.log 2>&1; echo "replay rc=$?" >> gpurun_out/r02_tests_replay.log
timeout 300 python scripts/measure_tf32_peak.py > gpurun_out/r02_tf32_peak.json 2> gpurun_out/r02_tf32_peak.err
timeout 600 python bench.py --steps 20 --warmup 5 --dump-profile gpurun_out/r02a_prof_resnet50.json > gpurun_out/r02a_bench_resnet50.json 2> gpurun_out/r02a_bench_resnet50.err
for m in vgg16 alexnet lenet; do timeout 600 python bench.py --model $m --steps 20 --warmup 5 --dump-profile gpurun_out/r02a_prof_$m.json > gpurun_out/r02a_bench_$m.json 2> gpurun_out/r02a_bench_$m.err; done
tail -3 gpurun_out/r02_tests_old.log; tail -5 gpurun_out/r02_tests_replay.log; cat gpurun_out/r02_tf32_peak.json; for m in resnet50 vgg16 alexnet lenet; do python -c "
import json,sys
try:
    j=json.load(open('gpurun_out/r02a_bench_$m.json')); print('$m', j['value'], j['ms_per_step'], j['e2e']['value'], j['roofline']['frac'], j.get('cpu_baseline',{}).get('value'))
except Exception as e: print('$m', 'FAILED', e)
"; done
