#!/bin/bash
# round 2, GPU call 3: precision variants (rRMSE + speed), wavefront v2b (compact paths, adaptive waves, batch refill)
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
echo "== wavefront quick check"; timeout 300 python tools/wf_check.py quick > $O/c3_wf_check.log 2>&1; echo "rc=$?"; grep -c "identical True" $O/c3_wf_check.log; grep "identical False" $O/c3_wf_check.log
echo "== FAST rRMSE by build"; timeout 900 python tools/fast_rrmse.py "default nostrict sfu nofmad precdiv allprec sfunofmad" 2 > $O/c3_fast_rrmse.log 2>&1; grep " FAST" $O/c3_fast_rrmse.log; grep -c "PARITY_FAST.*rRMSE 0.000e+00" $O/c3_fast_rrmse.log
echo "== A/B persistent (precision cost)"; timeout 900 python tools/ab.py "default nostrict sfu nofmad precdiv allprec sfunofmad" "is-1M-1080p:128 disney-1M-1080p:64" > $O/c3_ab_precision.log 2>&1; cat $O/c3_ab_precision.log
echo "== A/B wavefront"; timeout 900 python tools/ab.py "default wfc5 wfr2 wfr8 wfr16 wfh5 wfh3" "is-1M-1080p:128 disney-1M-1080p:512 lambert-100k-1080p:128" --pipeline 1 > $O/c3_ab_wavefront.log 2>&1; cat $O/c3_ab_wavefront.log
for wl in disney-1M-1080p:512 lambert-100k-1080p:128; do
  name=${wl%%:*}; nf=${wl##*:}
  echo "== launch list $name"; timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/c3_wf_launches_$name.csv python tools/wf_run.py $name 1 $nf > $O/c3_ncu_$name.log 2>&1; echo "rc=$?"
  python tools/launch_totals.py $O/c3_wf_launches_$name.csv
done
echo "== full capture wf_trace disney (bounce 0 of the second wave)"; timeout 900 ncu --set full --clock-control none --import-source on -k regex:wf_trace -s 5 -c 1 -f -o $O/c3_wf_trace_disney python tools/wf_run.py disney-1M-1080p 1 512 > $O/c3_ncu_full_disney.log 2>&1; echo "rc=$?"
ncu -i $O/c3_wf_trace_disney.ncu-rep --page raw --csv > $O/c3_wf_trace_disney_raw.csv 2>/dev/null
python tools/ncu_summary.py $O/c3_wf_trace_disney_raw.csv 2>/dev/null | grep -E "time_duration|thread_inst_executed_per|issue_active|inst_executed.sum |long_scoreboard|registers"
echo "== parity tests (pipelines)"; timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q -p no:cacheprovider --timeout 600 -k "pipeline or wavefront" > $O/c3_tests_parity.log 2>&1; echo "rc=$?"; tail -5 $O/c3_tests_parity.log
