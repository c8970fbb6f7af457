#!/bin/bash
# round 2, GPU call 35 (4 GPUs): the driver's scaling command at N = 4 on the final tree (pipelined combine, other workloads included)
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
N=4
echo "== bench --gpus $N"; SECONDS=0
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2962$N bench.py --gpus $N --steps 20 --warmup 5 > $O/c35_bench_$N.json 2> $O/c35_bench_$N.err; echo "rc=$? wall ${SECONDS}s"; grep -v "OMP_NUM_THREADS\|^\*\*\*" $O/c35_bench_$N.err | tail -3
python - <<PY
import json
try:
    d=[json.loads(l) for l in open('$O/c35_bench_$N.json') if l.startswith('{')][-1]
    for k in ('value','ms_per_step','e2e','parity','per_rank_kernel_ms','gpu_launches'): print(k, d.get(k))
    print({k:v['value'] for k,v in d['other_workloads'].items()})
except Exception as e: print('parse failed', e)
PY
