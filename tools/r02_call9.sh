#!/bin/bash
# round 2, GPU call 9: find the 3 % the persistent kernel lost against the round-1 build (same box)
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
for rep in 1 2; do
  echo -n "[r01] "; (cd _r01 && timeout 300 python tools/wf_run.py is-1M-1080p 2 256 2>&1 | tail -1)
done
timeout 900 python tools/ab.py "default tri64 node128 nofar n128nofar default" "is-1M-1080p:256 disney-1M-1080p:64 lambert-100k-1080p:64" --reps 2 2>&1 | tee $O/c9_ab.log
echo -n "[r01 lambert] "; (cd _r01 && timeout 300 python tools/wf_run.py lambert-100k-1080p 2 64 2>&1 | tail -1)
echo "== quick parity"; timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -p no:cacheprovider -k "oracle_10k or golden or pipeline_is_bit" 2>&1 | tail -2
