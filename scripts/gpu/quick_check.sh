timeout 600 python -m pytest tests/test_gpu_bn_maxpool.py -m gpu -x -q 2>&1 | tail -8
timeout 900 python -m pytest tests/test_gpu_network.py tests/test_gpu_replay.py -m gpu -x -q 2>&1 | tail -5
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('resnet50', round(j['value']), round(j['ms_per_step'],3), round(j['e2e']['value']), j['gpu_launches'])"
