#!/bin/bash
# round 2, GPU call 8: no-alloc entry point test, same-box A/B against the round-1 build, default bench + reference arm
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
echo "== no-alloc test alone (fresh process)"; timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > $O/c8_noalloc_alone.log 2>&1; echo "rc=$?"; tail -3 $O/c8_noalloc_alone.log
echo "== no-alloc test after other GPU tests"; timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_env_cdf.py tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > $O/c8_noalloc_after.log 2>&1; echo "rc=$?"; tail -3 $O/c8_noalloc_after.log
if [ -d _r01 ]; then
  echo "== r01 build vs now, same box, is-1M 256-frame launches (wf_run: second launch timed)"
  for rep in 1 2 3; do
    echo -n "[r01] "; (cd _r01 && timeout 300 python tools/wf_run.py is-1M-1080p 2 256 2>&1 | tail -1)
    echo -n "[now] "; timeout 300 python tools/wf_run.py is-1M-1080p 2 256 2>&1 | tail -1
  done
  echo -n "[r01 disney] "; (cd _r01 && timeout 300 python tools/wf_run.py disney-1M-1080p 2 64 2>&1 | tail -1)
  echo -n "[now disney] "; timeout 300 python tools/wf_run.py disney-1M-1080p 2 64 2>&1 | tail -1
fi
echo "== variants"; timeout 600 python tools/ab.py "default ld128 nostrict default ld128 nostrict" "is-1M-1080p:256" --reps 3 2>&1 | tee $O/c8_ab.log
echo "== default bench"; SECONDS=0; timeout 1500 python bench.py > $O/c8_bench.json 2> $O/c8_bench.err; echo "rc=$? wall ${SECONDS}s"; tail -2 $O/c8_bench.err; python - <<PY
import json
d=[json.loads(l) for l in open('$O/c8_bench.json') if l.startswith('{')][-1]
for k in ('value','ms_per_step','e2e','parity','e2e_interactive','gpu_launches','crop_like_reference','cpu_baseline'): print(k, d.get(k))
r=d['roofline']; print({k:r[k] for k in r if k not in ('hbm','note','l2')}); print(r.get('l2'))
print({k:v['value'] for k,v in d['other_workloads'].items()})
PY
echo "== reference arm"; SECONDS=0; timeout 900 python bench.py --impl reference > $O/c8_bench_ref.json 2>$O/c8_ref.err; echo "rc=$? wall ${SECONDS}s"; cut -c1-260 $O/c8_bench_ref.json
