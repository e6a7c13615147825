mkdir -p gpurun_out
for m in alexnet lenet; do
timeout 300 python scripts/timeline.py $m gpurun_out/timeline_$m.json 2>&1 | tail -1
python scripts/timeline_report.py gpurun_out/timeline_$m.json 3 > gpurun_out/timeline_report_$m.txt 2>/dev/null; head -40 gpurun_out/timeline_report_$m.txt | cut -c1-150
done
