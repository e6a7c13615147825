mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parallel.py -m gpu -q -x > gpurun_out/r02_tests_dp.log 2>&1; echo "dp test rc=$?"; tail -3 gpurun_out/r02_tests_dp.log
for n in 8 4 2; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r02_bench_n$n.json 2> gpurun_out/r02_bench_n$n.err
  echo "N=$n rc=$?"
  python -c "
import json
j=json.load(open('gpurun_out/r02_bench_n$n.json')); print('N=$n', round(j['value']), round(j['ms_per_step'],3), round(j['e2e']['value']), j.get('dp_check',{}).get('replicas_bit_identical'))
"
done
