#!/bin/bash
# round 2, GPU call 5: 256-bit loads, traversal-kernel stack variants, CDF-search sampler, bounds-check build
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
echo "== wavefront quick check"; timeout 300 python tools/wf_check.py quick > $O/c5_wf_check.log 2>&1; echo "rc=$?"; grep -c "identical True" $O/c5_wf_check.log; grep "identical False" $O/c5_wf_check.log
echo "== A/B persistent ld256"; timeout 600 python tools/ab.py "default ld128" "is-1M-1080p:128 disney-1M-1080p:64 lambert-100k-1080p:64 full-10M-4k:16" > $O/c5_ab_persistent.log 2>&1; cat $O/c5_ab_persistent.log
echo "== CDF-search sampler"; timeout 300 python tools/ab.py "default" "is-1M-1080p:128" --env 1 2>&1 | tail -1 | tee $O/c5_ab_cdf.log
echo "== A/B wavefront"; timeout 1200 python tools/ab.py "default ld128 wfc3 wfs8 wfs16 wfs24 wfc3s16 wfc3s32 wfc3r16" "disney-1M-1080p:512 lambert-100k-1080p:128" --pipeline 1 --reps 1 > $O/c5_ab_wavefront.log 2>&1; cat $O/c5_ab_wavefront.log
echo "== new tests"; timeout 1200 python -m pytest tests/test_env_cdf.py tests/test_gpu_parity.py -m gpu -q -p no:cacheprovider --timeout 600 -k "cdf or error or display or wavefront" > $O/c5_tests.log 2>&1; echo "rc=$?"; tail -6 $O/c5_tests.log
echo "== bounds-check build under the parity suite"; PT_LIBPT=$PWD/opengl_ray_tracing_b200/libpt_dbg.so timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_env_cdf.py -m gpu -q -p no:cacheprovider --timeout 600 > $O/c5_tests_dbg.log 2>&1; echo "rc=$?"; tail -4 $O/c5_tests_dbg.log
