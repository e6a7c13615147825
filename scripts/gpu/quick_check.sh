timeout 900 python -m pytest tests/test_gpu_helpers.py tests/test_gpu_network.py tests/test_gpu_replay.py -m gpu -x -q 2>&1 | tail -3
for m in resnet50 alexnet lenet; do
timeout 300 python bench.py --model $m --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$m', round(j['value']), round(j['ms_per_step'],3), round(j['e2e']['value']), j['gpu_launches'])"
done
