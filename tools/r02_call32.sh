#!/bin/bash
# round 2, GPU call 32: the metric kernel in its own translation unit with -ftz=true — A/B against the same tree without the flag, full GPU suite, bench
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 500 python tools/ab.py "default noftz default noftz" "is-1M-1080p:256" --reps 2 2>&1 | tee $O/c32_ab.log | tail -4
echo "== pytest -x -m gpu"; timeout 2400 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider > $O/c32_tests.log 2>&1; echo "rc=$?"; tail -3 $O/c32_tests.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
echo "== bench"; timeout 1500 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/c32_bench.json 2> $O/c32_bench.err; echo "rc=$?"; python - <<PY
import json
d=[json.loads(l) for l in open('$O/c32_bench.json') if l.startswith('{')][-1]
for k in ('value','ms_per_step','e2e','parity','gpu_launches','clocks'): print(k, d.get(k))
print(d['e2e_interactive'])
r=d['roofline']; print({k:r[k] for k in ('bound','achieved','peak','unit','frac','capture','no_issue_figure') if k in r})
print({k:v['value'] for k,v in d['other_workloads'].items()})
PY
