#!/bin/bash
# round 2, final check: the driver's GPU commands on the final tree
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
echo "== pytest -x -m gpu"; timeout 2400 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider > $O/c19_tests.log 2>&1; echo "rc=$?"; tail -3 $O/c19_tests.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
echo "== bench"; timeout 1500 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/c19_bench.json 2> $O/c19_bench.err; echo "rc=$?"; python - <<PY
import json
d=[json.loads(l) for l in open('$O/c19_bench.json') if l.startswith('{')][-1]
for k in ('value','ms_per_step','e2e','parity','gpu_launches','clocks'): print(k, d.get(k))
r=d['roofline']; print({k:r[k] for k in ('bound','achieved','peak','unit','frac','traffic','lanes','capture') if k in r}); print(r.get('l2')); print(r.get('fetches'))
print({k:v['value'] for k,v in d['other_workloads'].items()})
PY
