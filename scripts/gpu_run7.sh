mkdir -p gpurun_out
run() { timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/tmp_bench.json 2> gpurun_out/tmp_bench.err; python -c "
import json,sys
j=json.load(open('gpurun_out/tmp_bench.json')); print('$1', j['value'], j['ms_per_step'], j['e2e']['value'])
"; }
run default
DL4JB200_EAGER_REPACK=1 run eager_repack
DL4JB200_MAIN_PRIO=-1 run main_high
DL4JB200_MAIN_PRIO=-1 DL4JB200_EAGER_REPACK=1 run main_high_eager_repack
run default_again
