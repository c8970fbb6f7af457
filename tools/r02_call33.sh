#!/bin/bash
# round 2, GPU call 33 (2 GPUs): final tree — the 2-GPU tests (pt_combine synchronous / pipelined) and the driver's scaling command at N = 2
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
echo "== multi-GPU tests"; timeout 900 python -m pytest tests/test_multi_gpu.py tests/test_y_comm_single.py tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider --timeout 600 > $O/c33_tests.log 2>&1; echo "rc=$?"; tail -3 $O/c33_tests.log
N=2
echo "== bench --gpus $N"; SECONDS=0
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2963$N bench.py --gpus $N --steps 20 --warmup 5 > $O/c33_bench_$N.json 2> $O/c33_bench_$N.err; echo "rc=$? wall ${SECONDS}s"; grep -v "OMP_NUM_THREADS\|^\*\*\*" $O/c33_bench_$N.err | tail -3
python - <<PY
import json
try:
    d=[json.loads(l) for l in open('$O/c33_bench_$N.json') if l.startswith('{')][-1]
    for k in ('value','ms_per_step','e2e','parity','per_rank_kernel_ms','gpu_launches'): print(k, d.get(k))
    print({k:v['value'] for k,v in d['other_workloads'].items()})
except Exception as e: print('parse failed', e)
PY
echo "== reference arm under torchrun"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29641 bench.py --impl reference --gpus $N --steps 2 --warmup 1 2>/dev/null | cut -c1-300
