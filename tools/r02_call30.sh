#!/bin/bash
# round 2, GPU call 30: ncu --set full of the final persistent kernel on configs 3 and 2 (disney-1M, lambert-100k), 64-frame launches
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
for w in disney-1M-1080p lambert-100k-1080p; do
  echo "== $w un-profiled"; timeout 300 python tools/wf_run.py $w 2 64 2>&1 | tail -1
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_persistent -s 1 -c 1 -f -o $O/r02_$w python tools/wf_run.py $w 2 64 > $O/r02_${w}_ncu.log 2>&1; echo "ncu rc=$?"
  ncu -i $O/r02_$w.ncu-rep --page raw --csv > $O/r02_${w}_raw.csv 2>/dev/null
  python tools/ncu_summary.py $O/r02_${w}_raw.csv > $O/r02_${w}_summary.txt; wc -l $O/r02_${w}_summary.txt
  rm -f $O/r02_$w.ncu-rep
done
