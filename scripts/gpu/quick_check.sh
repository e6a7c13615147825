mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_conv_tc_gemm_tma_persist -s 3 -c 1 -o gpurun_out/r30_res2_3x3 python scripts/pair_bench.py "res2 3x3" > gpurun_out/r30_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/r30_ncu.log
