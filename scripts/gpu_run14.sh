mkdir -p gpurun_out
timeout 600 python scripts/timeline.py resnet50 gpurun_out/r02_timeline_resnet50.json 2>&1 | tail -1
python scripts/timeline_report.py gpurun_out/r02_timeline_resnet50.json > gpurun_out/r02_timeline_report.txt; head -6 gpurun_out/r02_timeline_report.txt
python scripts/tail_report.py gpurun_out/r02_timeline_resnet50.json 1600
