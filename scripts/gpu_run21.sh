mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_helpers.py tests/test_gpu_network.py tests/test_gpu_bwd_bnstats.py tests/test_zz_no_alloc.py -m gpu -q --maxfail=5 2>&1 | tail -3
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/tmp_bench.json 2> gpurun_out/tmp_bench.err
python -c "
import json
j=json.load(open('gpurun_out/tmp_bench.json')); print('resnet50', round(j['value']), round(j['ms_per_step'],3), round(j['e2e']['value']), round(j['roofline']['frac'],3), {k:(round(v['achieved_gbs']),round(v['share_of_step'],3)) for k,v in j['hbm_kernels'].items()})
"; tail -2 gpurun_out/tmp_bench.err
