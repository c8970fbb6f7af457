#!/bin/bash
# round 2, GPU call 12: final kernel — same-box yardstick, profiles (launch list + full capture), default bench, GPU suite
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
if [ -d _r01 ]; then echo -n "[r01] "; (cd _r01 && timeout 300 python tools/wf_run.py is-1M-1080p 2 256 2>&1 | tail -1); fi
echo -n "[now] "; timeout 300 python tools/wf_run.py is-1M-1080p 2 256 2>&1 | tail -1
timeout 600 python tools/ab.py "default" "disney-1M-1080p:64 lambert-100k-1080p:64 full-10M-4k:16" 2>&1 | tail -3
bash tools/profile_run.sh r02 2>&1 | tail -4
echo "== default bench"; SECONDS=0; timeout 1500 python bench.py > $O/c12_bench.json 2> $O/c12_bench.err; echo "rc=$? wall ${SECONDS}s"; tail -2 $O/c12_bench.err; python - <<PY
import json
d=[json.loads(l) for l in open('$O/c12_bench.json') if l.startswith('{')][-1]
for k in ('value','ms_per_step','e2e','parity','gpu_launches'): print(k, d.get(k))
r=d['roofline']; print({k:r[k] for k in r if k not in ('hbm','note','l2')})
print({k:v['value'] for k,v in d['other_workloads'].items()})
PY
echo "== all GPU tests"; timeout 2400 python -m pytest tests -x -m gpu -q -p no:cacheprovider --timeout 900 > $O/c12_tests.log 2>&1; echo "rc=$?"; tail -4 $O/c12_tests.log
