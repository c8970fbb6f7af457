#!/bin/bash
# round 2: the driver's GPU commands on the tree with the pipelined combine (libpt.so rebuilt: copy stream at high priority)
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
echo "== pytest -x -m gpu"; timeout 2400 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider > $O/c26_tests.log 2>&1; echo "rc=$?"; tail -3 $O/c26_tests.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
echo "== bench"; timeout 1500 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/c26_bench.json 2> $O/c26_bench.err; echo "rc=$?"; python - <<PY
import json
d=[json.loads(l) for l in open('$O/c26_bench.json') if l.startswith('{')][-1]
for k in ('value','ms_per_step','e2e','e2e_interactive','parity','gpu_launches','clocks'): print(k, d.get(k))
r=d['roofline']; print({k:r[k] for k in ('bound','achieved','peak','unit','frac','traffic','lanes','capture') if k in r})
print({k:v['value'] for k,v in d['other_workloads'].items()})
PY
echo "== reference arm"; timeout 900 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 | cut -c1-400
