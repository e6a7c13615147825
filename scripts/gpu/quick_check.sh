mkdir -p gpurun_out
timeout 600 python bench.py --model resnet50 --steps 20 --warmup 5 --dump-profile gpurun_out/r02_prof_resnet50.json > gpurun_out/r02_bench_resnet50.json 2> gpurun_out/r02_bench_resnet50.err
python -c "
import json
j=json.load(open('gpurun_out/r02_bench_resnet50.json')); print(round(j['value']), round(j['ms_per_step'],3), round(j['e2e']['value']), {k:(round(v['achieved_gbs']),round(v['frac_of_hbm'],3),round(v['share_of_step'],3)) for k,v in j['hbm_kernels'].items()}, j['cpu_baseline']['value'])"
tail -2 gpurun_out/r02_bench_resnet50.err
