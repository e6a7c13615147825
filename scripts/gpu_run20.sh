mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --maxfail=10 > gpurun_out/r02_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r02_tests.log
grep -n "^E  .*Error\|^E   *assert\|passed\|failed\|^FAILED" gpurun_out/r02_tests.log | cut -c1-260 | head -30
for m in vgg16 lenet alexnet resnet50; do
timeout 600 python bench.py --model $m --steps 20 --warmup 5 --no-cpu-baseline --dump-profile gpurun_out/tmp_prof_$m.json > gpurun_out/tmp_bench.json 2> gpurun_out/tmp_bench.err
python -c "
import json
j=json.load(open('gpurun_out/tmp_bench.json')); print('$m', round(j['value']), round(j['ms_per_step'],3), round(j['e2e']['value']), round(j['roofline']['frac'],3), round(j['roofline']['achieved'],1))
"; tail -2 gpurun_out/tmp_bench.err
done
