#!/bin/bash
# round 2, GPU call 7: final-build profiles (launch list + full capture), wf_trace capture, default bench, full GPU tests
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
bash tools/profile_run.sh r02 2>&1 | tail -15
echo "== wf_trace capture (final defaults), disney-1M, bounce 0 of the second wave"
timeout 300 python tools/wf_run.py disney-1M-1080p 1 512 2>&1 | tail -1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wf_trace -s 5 -c 1 -f -o $O/r02_wf_trace_disney python tools/wf_run.py disney-1M-1080p 1 512 > $O/r02_ncu_wf.log 2>&1; echo "rc=$?"
ncu -i $O/r02_wf_trace_disney.ncu-rep --page raw --csv > $O/r02_wf_trace_disney_raw.csv 2>/dev/null
python tools/ncu_summary.py $O/r02_wf_trace_disney_raw.csv > $O/r02_wf_trace_disney_summary.txt 2>/dev/null; grep -E "time_duration|thread_inst_executed_per|issue_active|l1tex__throughput|registers_per" $O/r02_wf_trace_disney_summary.txt
echo "== default bench"; /usr/bin/time -v -o $O/c7_bench.time timeout 1200 python bench.py > $O/c7_bench.json 2> $O/c7_bench.err; echo "rc=$?"; grep -E "Elapsed" $O/c7_bench.time; tail -2 $O/c7_bench.err; python - <<PY
import json
d=[json.loads(l) for l in open('$O/c7_bench.json') if l.startswith('{')][-1]
for k in ('value','ms_per_step','e2e','parity','e2e_interactive','gpu_launches'): print(k, d.get(k))
r=d['roofline']; print({k:r[k] for k in r if k not in ('hbm','note','l2')}); print(r.get('l2'))
print({k:v['value'] for k,v in d['other_workloads'].items()})
PY
echo "== reference arm"; /usr/bin/time -v -o $O/c7_ref.time timeout 900 python bench.py --impl reference > $O/c7_bench_ref.json 2>$O/c7_ref.err; echo "rc=$?"; grep Elapsed $O/c7_ref.time; cut -c1-200 $O/c7_bench_ref.json
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
echo "== all GPU tests"; timeout 2400 python -m pytest tests -m gpu -q -p no:cacheprovider --timeout 900 > $O/c7_tests.log 2>&1; echo "rc=$?"; tail -4 $O/c7_tests.log
