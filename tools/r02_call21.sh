#!/bin/bash
# round 2: bench.py after the interactive loop moved to a pinned display buffer (bench.py only; kernels unchanged)
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
echo "== bench"; timeout 1500 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/c21_bench.json 2> $O/c21_bench.err; echo "rc=$?"; python - <<PY
import json
d=[json.loads(l) for l in open('$O/c21_bench.json') if l.startswith('{')][-1]
for k in ('value','ms_per_step','e2e','e2e_interactive','parity','gpu_launches','basic_gpu'): print(k, d.get(k))
r=d['roofline']; print({k:r[k] for k in ('bound','achieved','peak','unit','frac','capture') if k in r})
PY
echo "== reference arm"; timeout 900 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 | cut -c1-600
