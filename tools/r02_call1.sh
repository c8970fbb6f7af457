#!/bin/bash
# round 2, GPU call 1: new config parity tests, wavefront v2 sanity, L2 probe, A/B of stack/fold variants, sanitizer on smoke()
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $O/c1_gpu.txt; nproc >> $O/c1_gpu.txt
echo "== wavefront quick check"; timeout 300 python tools/wf_check.py quick > $O/c1_wf_check.log 2>&1; echo "rc=$?"; cat $O/c1_wf_check.log | tail -30
echo "== config tests"; timeout 1500 python -m pytest tests/test_gpu_configs.py -m gpu -q -p no:cacheprovider --timeout 900 > $O/c1_tests_configs.log 2>&1; echo "rc=$?"; tail -15 $O/c1_tests_configs.log
echo "== L2 probe"; timeout 120 python -c "
from opengl_ray_tracing_b200 import renderer as R
for w in (16, 32, 48, 64, 96):
    print('window %d MiB: L2 %.0f GB/s, HBM(2 GiB) %.0f GB/s' % ((w,) + R.measure_l2_read_gbs(0, w, 64)))" > $O/c1_l2.log 2>&1; cat $O/c1_l2.log
echo "== A/B persistent"; timeout 900 python tools/ab.py "default foldv4 smem8 smem16 smem24 smem40 tos smem16tos c4smem32 c4smem48" "is-1M-1080p:128 disney-1M-1080p:64 lambert-100k-1080p:64" > $O/c1_ab_persistent.log 2>&1; cat $O/c1_ab_persistent.log
echo "== A/B wavefront"; timeout 900 python tools/ab.py "default wfc4 wfc6 wfr1 wfr8 wfr16 wfc4s16 wfc4r8" "is-1M-1080p:128 disney-1M-1080p:64 lambert-100k-1080p:64" --pipeline 1 > $O/c1_ab_wavefront.log 2>&1; cat $O/c1_ab_wavefront.log
echo "== parity tests"; timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q -p no:cacheprovider --timeout 600 > $O/c1_tests_parity.log 2>&1; echo "rc=$?"; tail -15 $O/c1_tests_parity.log
echo "== memcheck smoke"; timeout 600 compute-sanitizer --tool memcheck --log-file $O/c1_memcheck_smoke.log python __graft_entry__.py smoke > $O/c1_memcheck_smoke.out 2>&1; echo "rc=$?"; tail -3 $O/c1_memcheck_smoke.log; tail -2 $O/c1_memcheck_smoke.out
