mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_bwd_bnstats.py tests/test_gpu_network.py -m gpu -q 2>&1 | tail -2
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/tmp_bench.json 2> gpurun_out/tmp_bench.err
python -c "
import json,sys
j=json.load(open('gpurun_out/tmp_bench.json')); print('resnet50', j['value'], j['ms_per_step'], j['e2e']['value'], j['roofline']['frac'], j['roofline']['achieved'])
"
