#!/bin/bash
# round 2, GPU call 6 (2 GPUs): pt_comm / pt_combine on hardware — test, ptdrive --gpus 2, bench --gpus 2
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -4
echo "== multi-GPU test"; timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -q -p no:cacheprovider --timeout 500 > $O/c6_tests.log 2>&1; echo "rc=$?"; tail -15 $O/c6_tests.log
D=opengl_ray_tracing_b200
echo "== ptdrive 1 GPU / 2 GPUs"
timeout 300 $D/ptdrive --variant 2 --icosphere 71 --size 1920x1080 --spp 64 --out $O/c6_one.png --pfm $O/c6_one.pfm 2>&1 | tail -3
timeout 300 $D/ptdrive --variant 2 --icosphere 71 --size 1920x1080 --spp 64 --gpus 2 --split tile --out $O/c6_tile.png --pfm $O/c6_tile.pfm 2>&1 | tail -3
timeout 300 $D/ptdrive --variant 2 --icosphere 71 --size 1920x1080 --spp 64 --gpus 2 --split sample --out $O/c6_sample.png 2>&1 | tail -3
cmp $O/c6_one.png $O/c6_tile.png && echo "PNG tile == one"; cmp $O/c6_one.pfm $O/c6_tile.pfm && echo "PFM tile == one (bit-identical)"
rm -f $O/c6_*.pfm
for split in sample tile; do
  echo "== bench --gpus 2 --split $split"
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --split $split $( [ $split = tile ] && echo --no-other-workloads ) > $O/c6_bench_$split.json 2> $O/c6_bench_$split.err; echo "rc=$?"; tail -3 $O/c6_bench_$split.err
  python - <<PY
import json
try:
    d=[json.loads(l) for l in open('$O/c6_bench_$split.json') if l.startswith('{')][-1]
    for k in ('value','ms_per_step','e2e','parity','per_rank_kernel_ms','other_workloads'): print(k, d.get(k))
    print(d['config'].get('combine'))
except Exception as e: print('parse failed', e)
PY
done
