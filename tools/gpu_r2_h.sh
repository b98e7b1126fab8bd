#!/bin/bash
# round 2, eighth GPU pass: the super-sphere level -- checked build under the tests, the whole GPU suite,
# A/B against the build without it, and every rank's share of an 8-way split over tile heights.
set -u
mkdir -p gpurun_out
RT_B200_TEST_LIB=build/variants/checked/librt_b200.so timeout 600 python -m pytest tests/test_gpu_sweep.py tests/test_gpu_forest.py tests/test_gpu_grazing.py tests/test_gpu_parity.py -x -q \
   -k "not true_size and not true_resolution" > gpurun_out/pytest_checked.log 2>&1; echo "pytest checked rc=$?"; tail -n 3 gpurun_out/pytest_checked.log | cut -c1-300
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/pytest_gpu.log | cut -c1-300
FT="timeout 300 python tools/frame_time.py"
( $FT c4 reps=4 -- world=8 reps=4 -- want_heavy=1 -- want_heavy=4 -- world=2 -- world=4
  for v in nosupers bulk1; do $FT c4 lib=build/variants/$v/librt_b200.so reps=4 -- world=8 reps=4; done
  $FT c3; $FT c3 lib=build/variants/nosupers/librt_b200.so; $FT c2; $FT c2 lib=build/variants/nosupers/librt_b200.so ) > gpurun_out/ab8.log 2>&1
echo "ab rc=$?"; cat gpurun_out/ab8.log
timeout 600 python tools/explore_tile.py 8 > gpurun_out/explore_tile_w8.log 2>&1; cat gpurun_out/explore_tile_w8.log
