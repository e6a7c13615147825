mkdir -p gpurun_out
timeout 600 python scripts/timeline.py resnet50 gpurun_out/timeline_r26.json 2>&1 | tail -2
python scripts/timeline_report.py gpurun_out/timeline_r26.json 3 | head -12
