#!/bin/bash
# One profiling gpurun call (tools/make_profiles.py turns its files into profiles/): tools/profile_run.sh <tag>
# Order: the un-profiled command first; then the launch list; then ONE --set full capture of the render kernel.
cd "$(dirname "$0")/.."
tag=${1:-r01}; O=gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-other-workloads"
$CMD > $O/${tag}_bench_short.json 2> $O/${tag}_bench_short.err || { echo "bench failed"; tail $O/${tag}_bench_short.err; exit 1; }
tail -c 600 $O/${tag}_bench_short.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${tag}_launches.csv $CMD > $O/${tag}_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_persistent -s 3 -c 1 -f -o $O/${tag}_render_persistent $CMD > $O/${tag}_ncu_full.log 2>&1
ncu -i $O/${tag}_render_persistent.ncu-rep --page raw --csv > $O/${tag}_render_persistent_raw.csv 2>/dev/null
ncu -i $O/${tag}_render_persistent.ncu-rep --page source --csv --print-source cuda,sass > $O/${tag}_src_cs.csv 2>/dev/null
ls -la $O | tail -12
