#!/bin/bash
# round 2, GPU call 15 (2 GPUs): PT_COMBINE_ASYNC test, bench --gpus 2 with overlapped copies
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
echo "== multi-GPU test"; timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q -p no:cacheprovider --timeout 500 > $O/c15_tests.log 2>&1; echo "rc=$?"; tail -5 $O/c15_tests.log
echo "== bench --gpus 2"; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 10 --warmup 3 --no-other-workloads > $O/c15_bench.json 2> $O/c15_bench.err; echo "rc=$?"; grep -v "OMP_NUM\|^\*\*" $O/c15_bench.err | tail -3
python - <<PY
import json
try:
    d=[json.loads(l) for l in open('$O/c15_bench.json') if l.startswith('{')][-1]
    for k in ('value','ms_per_step','e2e','parity','per_rank_kernel_ms'): print(k, d.get(k))
except Exception as e: print('parse failed', e)
PY
