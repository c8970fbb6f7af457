#!/bin/bash
# A/B of kernel builds on the GPU box (tools/build_variant.py): usage tools/ab_run.sh "<variants>" "<workload:frames> ..."
# "default" = the in-tree libpt.so.
cd "$(dirname "$0")/.."
for wl in $2; do
  name=${wl%%:*}; nf=${wl##*:}
  for v in $1; do
    if [ "$v" = default ]; then lib=""; else lib="$PWD/opengl_ray_tracing_b200/libpt_$v.so"; fi
    for rep in 1 2; do
      echo -n "[$v] "; PT_LIBPT=$lib timeout 300 python tools/wf_run.py $name 2 $nf 2>&1 | tail -1
    done
  done
done
