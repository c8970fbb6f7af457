#!/bin/bash
# round 2, GPU call 14: tile-order table A/B, async readback test, bench with overlapped copies
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
if [ -d _r01 ]; then echo -n "[r01] "; (cd _r01 && timeout 300 python tools/wf_run.py is-1M-1080p 2 256 2>&1 | tail -1); fi
timeout 900 python tools/ab.py "default notab default notab" "is-1M-1080p:256 disney-1M-1080p:64 lambert-100k-1080p:64" --reps 2 2>&1 | tee $O/c14_ab.log
echo "== tests"; timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider --timeout 600 > $O/c14_tests.log 2>&1; echo "rc=$?"; tail -4 $O/c14_tests.log
echo "== bench"; timeout 900 python bench.py --no-other-workloads > $O/c14_bench.json 2> $O/c14_bench.err; echo "rc=$?"; tail -2 $O/c14_bench.err; python - <<PY
import json
d=[json.loads(l) for l in open('$O/c14_bench.json') if l.startswith('{')][-1]
for k in ('value','ms_per_step','e2e','parity','gpu_launches'): print(k, d.get(k))
PY
