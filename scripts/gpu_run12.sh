mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_bwd_bnstats.py -m gpu -q 2>&1 | tail -2
timeout 600 python scripts/timeline.py resnet50 gpurun_out/r02_timeline_resnet50.json 2>&1 | tail -1
python scripts/timeline_report.py gpurun_out/r02_timeline_resnet50.json > gpurun_out/r02_timeline_report.txt; head -45 gpurun_out/r02_timeline_report.txt
