#!/bin/bash
# round 2, GPU call 22 (2 GPUs): PT_COMBINE_PIPELINED on hardware — tests, then bench --gpus 2 with the pipelined and the synchronous combine
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -4
echo "== multi-GPU tests"; timeout 900 python -m pytest tests/test_multi_gpu.py tests/test_y_comm_single.py -m gpu -q -p no:cacheprovider --timeout 600 > $O/c22_tests.log 2>&1; echo "rc=$?"; tail -15 $O/c22_tests.log
for how in pipelined sync; do
  echo "== bench --gpus 2 --combine $how"
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 2 --steps 10 --warmup 3 --combine $how --no-other-workloads > $O/c22_bench_$how.json 2> $O/c22_bench_$how.err; echo "rc=$?"; grep -v "OMP_NUM_THREADS\|^\*\*\*" $O/c22_bench_$how.err | tail -3
  python - <<PY
import json
try:
    d=[json.loads(l) for l in open('$O/c22_bench_$how.json') if l.startswith('{')][-1]
    for k in ('value','ms_per_step','e2e','parity','per_rank_kernel_ms','render_stream_ms_per_step_rank0'): print(k, d.get(k))
    print(d['config'].get('combine'))
except Exception as e: print('parse failed', e)
PY
done
