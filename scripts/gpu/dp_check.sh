mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parallel.py -m gpu -q -x 2>&1 | tail -3
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02_bench_n2.json 2> gpurun_out/r02_bench_n2.err
python -c "
import json
j=json.load(open('gpurun_out/r02_bench_n2.json')); print('N=2', round(j['value']), round(j['ms_per_step'],3), round(j['e2e']['value']), j.get('dp_check',{}).get('replicas_bit_identical'))"
