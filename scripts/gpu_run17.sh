mkdir -p gpurun_out
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
j=json.load(open('gpurun_out/r02_bench_reference_$m.json')); print('ref $m', j['value'], j['ms_per_step'], j['cpu_baseline']['cores'], j['cpu_baseline']['threads'])
"
done
OMP_NUM_THREADS=1 timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_reference_resnet50_omp1env.json 2>/dev/null; python -c "
import json
j=json.load(open('gpurun_out/r02_bench_reference_resnet50_omp1env.json')); print('ref resnet50 with OMP_NUM_THREADS=1 in the environment (torchrun default)', j['value'])"
