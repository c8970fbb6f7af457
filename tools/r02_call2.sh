#!/bin/bash
# round 2, GPU call 2: config-4/5 tests after the direction-precision fix, per-kernel times of the wavefront pipeline, captures
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
echo "== config tests"; timeout 1500 python -m pytest tests/test_gpu_configs.py -m gpu -q -p no:cacheprovider --timeout 900 -k "config4 or config5 or config3" > $O/c2_tests_configs.log 2>&1; echo "rc=$?"; tail -8 $O/c2_tests_configs.log
echo "== bench short"; timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > $O/c2_bench.json 2> $O/c2_bench.err; echo "rc=$?"; python -c "
import json; d=json.load(open('$O/c2_bench.json')); print(d['value'], d['ms_per_step'], d['e2e'])"
for wl in disney-1M-1080p is-1M-1080p lambert-100k-1080p; do
  echo "== plain $wl"; timeout 300 python tools/wf_run.py $wl 1 64 2>&1 | tail -1
  echo "== launch list $wl"; timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/c2_wf_launches_$wl.csv python tools/wf_run.py $wl 1 64 > $O/c2_ncu_$wl.log 2>&1; echo "rc=$?"
  python - <<PY
import csv, collections
rows=[r for r in csv.reader(open('$O/c2_wf_launches_$wl.csv')) if len(r)>5]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value'); ui=hdr.index('Metric Unit')
agg=collections.OrderedDict()
for r in rows[1:]:
    try: v=float(r[vi].replace(',',''))
    except ValueError: continue
    u=r[ui]; v=v/1e3 if u in ('usecond','us') else v/1e6 if u in ('nsecond','ns') else v
    k=r[ki][:60]; a=agg.setdefault(k,[0,0.0]); a[0]+=1; a[1]+=v
for k,(n,t) in agg.items(): print('%-62s n=%3d total %.3f ms'%(k,n,t))
PY
done
echo "== full capture wf_trace disney"; timeout 900 ncu --set full --clock-control none --import-source on -k regex:wf_trace -s 1 -c 1 -f -o $O/c2_wf_trace_disney python tools/wf_run.py disney-1M-1080p 1 64 > $O/c2_ncu_full_disney.log 2>&1; echo "rc=$?"
ncu -i $O/c2_wf_trace_disney.ncu-rep --page raw --csv > $O/c2_wf_trace_disney_raw.csv 2>/dev/null
echo "== full capture wf_trace is-1M"; timeout 900 ncu --set full --clock-control none --import-source on -k regex:wf_trace -s 1 -c 1 -f -o $O/c2_wf_trace_is python tools/wf_run.py is-1M-1080p 1 64 > $O/c2_ncu_full_is.log 2>&1; echo "rc=$?"
ncu -i $O/c2_wf_trace_is.ncu-rep --page raw --csv > $O/c2_wf_trace_is_raw.csv 2>/dev/null
ls -la $O | tail -20
