#!/bin/bash
# round 2, GPU call 11: which of the three kernel-side changes of the first round-2 commit costs the persistent kernel 2.3 %
cd "$(dirname "$0")/.."
echo -n "[r01] "; (cd _r01 && timeout 300 python tools/wf_run.py is-1M-1080p 2 256 2>&1 | tail -1)
timeout 900 python tools/ab.py "default plain plainnofar plainv4 plainall default plain plainnofar plainv4 plainall" "is-1M-1080p:256" --reps 2 2>&1 | tee gpurun_out/c11_ab.log
echo -n "[r01] "; (cd _r01 && timeout 300 python tools/wf_run.py is-1M-1080p 2 256 2>&1 | tail -1)
