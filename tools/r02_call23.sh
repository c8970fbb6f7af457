#!/bin/bash
# round 2, GPU call 23 (8 GPUs): bench --gpus 8 with the pipelined and the synchronous combine, same box
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
for how in pipelined sync; do
  echo "== bench --gpus 8 --combine $how"; SECONDS=0
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus 8 --steps 20 --warmup 5 --combine $how --no-other-workloads > $O/c23_bench_8_$how.json 2> $O/c23_bench_8_$how.err; echo "rc=$? wall ${SECONDS}s"; grep -v "OMP_NUM_THREADS\|^\*\*\*" $O/c23_bench_8_$how.err | tail -3
  python - <<PY
import json
try:
    d=[json.loads(l) for l in open('$O/c23_bench_8_$how.json') if l.startswith('{')][-1]
    for k in ('value','ms_per_step','e2e','parity','per_rank_kernel_ms','render_stream_ms_per_step_rank0'): print(k, d.get(k))
except Exception as e: print('parse failed', e)
PY
done
