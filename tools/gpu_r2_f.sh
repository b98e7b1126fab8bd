#!/bin/bash
# round 2, sixth GPU pass: the checked build (device-side bounds checks; compute-sanitizer is closed on this
# pool) under the sweep / forest / parity tests, A/B of prologue preload, pipelined exact stage, brute unroll 1
# + thread-copied table, and the want_heavy knob.
set -u
mkdir -p gpurun_out
RT_B200_TEST_LIB=build/variants/checked/librt_b200.so timeout 900 python -m pytest tests/test_gpu_sweep.py tests/test_gpu_forest.py tests/test_gpu_grazing.py tests/test_gpu_parity.py -x -q \
   -k "not true_size and not true_resolution" > gpurun_out/pytest_checked.log 2>&1; echo "pytest checked rc=$?"; tail -n 5 gpurun_out/pytest_checked.log
grep -c "RT_CHECK failed" gpurun_out/pytest_checked.log
FT="timeout 300 python tools/frame_time.py c4"
( $FT reps=4 -- world=8 reps=4 -- want_heavy=8 -- want_heavy=12 -- want_heavy=16 -- want_heavy=24 -- want_heavy=32 -- world=8 want_heavy=16 -- world=8 want_heavy=24 -- world=4 want_heavy=24 -- world=4
  for v in nopreload pipe64 pipe64s32 brute1sync; do $FT lib=build/variants/$v/librt_b200.so reps=4 -- world=8 reps=4; done ) > gpurun_out/ab6_c4.log 2>&1
echo "ab rc=$?"; cat gpurun_out/ab6_c4.log
ls -la gpurun_out
