mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests/test_gpu_parallel.py -m gpu -q -x > gpurun_out/r02_tests_dp.log 2>&1; echo "rc=$?" >> gpurun_out/r02_tests_dp.log
tail -30 gpurun_out/r02_tests_dp.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02_bench_n2.json 2> gpurun_out/r02_bench_n2.err
tail -5 gpurun_out/r02_bench_n2.err
python -c "
import json
j=json.load(open('gpurun_out/r02_bench_n2.json')); print('N=2', j['value'], j['ms_per_step'], j['e2e']['value'], j.get('dp_check'))
"
