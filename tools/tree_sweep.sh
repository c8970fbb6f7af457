#!/bin/bash
# sweep of the production tree's build parameters (env, read at upload): tools/tree_sweep.sh "<workload:frames> ..."
cd "$(dirname "$0")/.."
for wl in $1; do
  name=${wl%%:*}; nf=${wl##*:}
  for cfg in "4 100" "2 100" "3 100" "6 100" "8 100" "4 50" "4 200" "8 200"; do
    set -- $cfg
    echo -n "[leaf_max $1 cnode $2%] "; PT_LEAF_MAX=$1 PT_SAH_CNODE_PCT=$2 timeout 300 python tools/wf_run.py $name 2 $nf 2>&1 | tail -1
  done
done
