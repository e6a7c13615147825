mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_helpers.py tests/test_gpu_bn_maxpool.py tests/test_gpu_network.py -m gpu -x -q -k "pool or subsampling or maxpool or resnet or lenet or alexnet or vgg" 2>&1 | tail -3
for m in resnet50 alexnet; do
timeout 300 python bench.py --model $m --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$m', round(j['value']), round(j['ms_per_step'],3), round(j['e2e']['value']), round(j['roofline']['frac'],3), {k:(round(v['achieved_gbs']),round(v['share_of_step'],3)) for k,v in j.get('hbm_kernels',{}).items()})"
done
