#!/bin/bash
# round 2, GPU call 17: ray-queue sorting in the wavefront pipeline (PT_WF_SORT) — bit-identity, A/B, launch list
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
V=$PWD/opengl_ray_tracing_b200/libpt_wfsort.so
echo "== bit-identity of the sorted pipeline"; PT_LIBPT=$V timeout 300 python tools/wf_check.py quick > $O/c17_wf_check.log 2>&1; echo "rc=$?"; grep -c "identical True" $O/c17_wf_check.log; grep "identical False" $O/c17_wf_check.log
PT_LIBPT=$V timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -p no:cacheprovider -k "wavefront or pipeline" 2>&1 | tail -2
echo "== A/B wavefront"; timeout 1200 python tools/ab.py "default wfsort" "disney-1M-1080p:512 lambert-100k-1080p:128 is-1M-1080p:128 full-10M-4k:16" --pipeline 1 --reps 2 2>&1 | tee $O/c17_ab.log
echo "== persistent"; timeout 600 python tools/ab.py "default" "disney-1M-1080p:512" --reps 1 2>&1 | tail -1
echo "== launch list disney (sorted)"; PT_LIBPT=$V timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/c17_wf_launches.csv python tools/wf_run.py disney-1M-1080p 1 512 > $O/c17_ncu.log 2>&1; echo "rc=$?"
python tools/launch_totals.py $O/c17_wf_launches.csv | grep -v "k_\|cub::DeviceScan\|material_derive"
