#!/bin/bash
# round 2, GPU call 16: final capture (launch list + full) of the final kernels, full GPU suite, default bench + reference arm
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
bash tools/profile_run.sh r02 2>&1 | tail -3
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
echo "== all GPU tests"; timeout 2400 python -m pytest tests -x -m gpu -q -p no:cacheprovider --timeout 900 > $O/c16_tests.log 2>&1; echo "rc=$?"; tail -4 $O/c16_tests.log
echo "== reference arm"; SECONDS=0; timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > $O/c16_bench_ref.json 2>$O/c16_ref.err; echo "rc=$? wall ${SECONDS}s"; cut -c1-200 $O/c16_bench_ref.json
echo "== default bench (driver's arguments)"; SECONDS=0; timeout 1500 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/c16_bench.json 2> $O/c16_bench.err; echo "rc=$? wall ${SECONDS}s"; tail -2 $O/c16_bench.err; python - <<PY
import json
d=[json.loads(l) for l in open('$O/c16_bench.json') if l.startswith('{')][-1]
for k in ('value','ms_per_step','e2e','parity','gpu_launches','crop_like_reference','e2e_interactive','cpu_baseline'): print(k, d.get(k))
r=d['roofline']; print({k:r[k] for k in r if k not in ('hbm','note','l2')}); print(r.get('l2'))
print({k:v['value'] for k,v in d['other_workloads'].items()})
PY
