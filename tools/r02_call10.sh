#!/bin/bash
# round 2, GPU call 10: bisect the 3 % regression of the persistent kernel over this round's commits (same box)
cd "$(dirname "$0")/.."
for rep in 1 2; do
  for d in _r01 _b1 _b2 _b3 _b4 .; do
    [ -d $d/opengl_ray_tracing_b200 ] || continue
    echo -n "[$d] "; (cd $d && timeout 300 python tools/wf_run.py is-1M-1080p 2 256 2>&1 | tail -1)
  done
done
