mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_conv_tc_gemm_tma_persist -s 27 -c 1 -o gpurun_out/r29_res3_3x3_pair python scripts/pair_bench.py "res3 3x3" > gpurun_out/r29_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/r29_ncu.log
