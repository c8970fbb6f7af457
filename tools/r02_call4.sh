#!/bin/bash
# round 2, GPU call 4: vote-scheduled traversal kernel, new bench.py, full GPU test suite
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
echo "== wavefront quick check"; timeout 300 python tools/wf_check.py quick > $O/c4_wf_check.log 2>&1; echo "rc=$?"; grep -c "identical True" $O/c4_wf_check.log; grep "identical False" $O/c4_wf_check.log
echo "== A/B wavefront"; timeout 900 python tools/ab.py "default wfc3 wfc5 wfr8 wfr16 wfnb8 wfnbm8 wfr8c3 wfs16" "disney-1M-1080p:512 lambert-100k-1080p:128 is-1M-1080p:128" --pipeline 1 > $O/c4_ab_wavefront.log 2>&1; cat $O/c4_ab_wavefront.log
echo "== persistent reference points"; timeout 300 python tools/ab.py "default" "disney-1M-1080p:512 lambert-100k-1080p:128 is-1M-1080p:128 full-10M-4k:16" > $O/c4_ab_persistent.log 2>&1; cat $O/c4_ab_persistent.log
echo "== wavefront full-10M-4k"; timeout 300 python tools/ab.py "default" "full-10M-4k:16" --pipeline 1 2>&1 | tail -1
for wl in disney-1M-1080p:512; do
  name=${wl%%:*}; nf=${wl##*:}
  echo "== launch list $name"; timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/c4_wf_launches_$name.csv python tools/wf_run.py $name 1 $nf > $O/c4_ncu_$name.log 2>&1; echo "rc=$?"
  python tools/launch_totals.py $O/c4_wf_launches_$name.csv | grep "pt::"
done
echo "== full capture wf_trace disney (bounce 0 of the second wave)"; timeout 900 ncu --set full --clock-control none --import-source on -k regex:wf_trace -s 5 -c 1 -f -o $O/c4_wf_trace_disney python tools/wf_run.py disney-1M-1080p 1 512 > $O/c4_ncu_full_disney.log 2>&1; echo "rc=$?"
ncu -i $O/c4_wf_trace_disney.ncu-rep --page raw --csv > $O/c4_wf_trace_disney_raw.csv 2>/dev/null
python tools/ncu_summary.py $O/c4_wf_trace_disney_raw.csv 2>/dev/null | grep -E "time_duration|thread_inst_executed_per|issue_active|inst_executed.sum |long_scoreboard|registers|l1tex__t_sector_hit|local_"
echo "== bench (new)"; timeout 900 python bench.py --steps 5 --warmup 3 > $O/c4_bench.json 2> $O/c4_bench.err; echo "rc=$?"; tail -3 $O/c4_bench.err; python - <<PY
import json
try:
    d=json.load(open('$O/c4_bench.json'))
    for k in ('value','ms_per_step','e2e','parity','crop_like_reference','e2e_interactive','other_workloads','cpu_baseline'): print(k, d.get(k))
    r=d['roofline']; print('roofline', {k:r[k] for k in r if k not in ('hbm','note')})
except Exception as e: print('bench parse failed', e); print(open('$O/c4_bench.json').read()[:2000])
PY
echo "== reference arm"; timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/c4_bench_ref.json 2> $O/c4_bench_ref.err; echo "rc=$?"; cat $O/c4_bench_ref.json | cut -c1-600
echo "== all GPU tests"; timeout 2400 python -m pytest tests -m gpu -q -p no:cacheprovider --timeout 900 > $O/c4_tests.log 2>&1; echo "rc=$?"; tail -8 $O/c4_tests.log
