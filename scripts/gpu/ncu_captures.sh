mkdir -p gpurun_out
python scripts/one_step.py 1 1 > gpurun_out/one_step_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r02_launches_all.csv python scripts/one_step.py 1 1 > gpurun_out/ncu1.log 2>&1
echo "launch list rc=$?"
for spec in "wgrad:k_conv_tc_wgrad_tma:30" "finalize:k_wgrad_finalize:30" "bnpartial:k_bn_partial_nhwc:4" "bnapply:k_bn_apply_nhwc:60" "bstgemm:k_conv_tc_gemm_tma_persist:150"; do
  name=${spec%%:*}; rest=${spec#*:}; pat=${rest%%:*}; skip=${rest##*:}
  ncu --set full --clock-control none --import-source on -k regex:$pat -s $skip -c 2 -o gpurun_out/r02_full_$name python scripts/one_step.py 1 1 > gpurun_out/ncu_$name.log 2>&1
  echo "$name rc=$?"
done
ls -la gpurun_out/*.ncu-rep
