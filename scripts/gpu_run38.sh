python - <<'PY'
import sys; sys.path.insert(0,'.')
import numpy as np, torch
from deeplearning4j_b200 import network, zoo
m = zoo.alexnet()
net = network.Network(m, 128)
print("stem:", net.stem)
PY
timeout 300 python bench.py --model alexnet --steps 20 --warmup 5 --no-cpu-baseline --dump-profile gpurun_out/prof_alex.json > /dev/null 2>&1
python - <<'PY'
import json
d=json.load(open('gpurun_out/prof_alex.json'))['per_layer']
for k,v in sorted(d.items(), key=lambda kv:-kv[1]['ms'])[:8]: print(k, round(v['ms'],4), round(v.get('tflops',0) or 0,1))
PY
