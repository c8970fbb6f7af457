#!/bin/bash
# round 2, GPU call 13 (8 GPUs): the driver's scaling command at N = 8 (and 4), ptdrive --gpus 8
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv,noheader | wc -l
for N in 8 4; do
  echo "== bench --gpus $N (sample split, default)"; SECONDS=0
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 10 --warmup 3 > $O/c13_bench_$N.json 2> $O/c13_bench_$N.err; echo "rc=$? wall ${SECONDS}s"; grep -v "OMP_NUM_THREADS\|^\*\*\*" $O/c13_bench_$N.err | tail -3
  python - <<PY
import json
try:
    d=[json.loads(l) for l in open('$O/c13_bench_$N.json') if l.startswith('{')][-1]
    for k in ('value','ms_per_step','e2e','parity','per_rank_kernel_ms','other_workloads'): print(k, d.get(k))
    print(d['config'].get('combine'))
except Exception as e: print('parse failed', e)
PY
done
echo "== bench --gpus 8 --split tile"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 8 --steps 10 --warmup 3 --split tile --no-other-workloads > $O/c13_bench_8_tile.json 2> $O/c13_bench_8_tile.err; echo "rc=$?"
python - <<PY
import json
try:
    d=[json.loads(l) for l in open('$O/c13_bench_8_tile.json') if l.startswith('{')][-1]
    for k in ('value','ms_per_step','e2e','parity','per_rank_kernel_ms'): print(k, d.get(k))
except Exception as e: print('parse failed', e)
PY
echo "== reference arm under torchrun"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --impl reference --gpus 8 --steps 3 --warmup 1 2>/dev/null | grep '^{' | cut -c1-300
echo "== ptdrive --gpus 8"; timeout 300 opengl_ray_tracing_b200/ptdrive --variant 2 --icosphere 71 --size 1920x1080 --spp 256 --gpus 8 --out $O/c13_8.png 2>&1 | tail -2
timeout 300 opengl_ray_tracing_b200/ptdrive --variant 2 --icosphere 71 --size 1920x1080 --spp 256 --out $O/c13_1.png 2>&1 | tail -2
cmp $O/c13_1.png $O/c13_8.png && echo "PNG 8 GPUs (tile split) == 1 GPU"
