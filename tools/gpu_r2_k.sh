#!/bin/bash
# round 2, eleventh GPU pass: work sharing inside a sweep launch (idle warps take half of a busy warp's unswept
# slots) -- checked build and the whole suite first, then share on / off at 1 .. 1/8 of the frame.
set -u
mkdir -p gpurun_out
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/smoke.log
RT_B200_TEST_LIB=build/variants/checked/librt_b200.so timeout 600 python -m pytest tests/test_gpu_sweep.py tests/test_gpu_forest.py tests/test_gpu_grazing.py tests/test_gpu_parity.py -x -q \
   -k "not true_size and not true_resolution" > gpurun_out/pytest_checked.log 2>&1; echo "pytest checked rc=$?"; tail -n 3 gpurun_out/pytest_checked.log | cut -c1-300
grep -c "shared work did not end\|RT_CHECK failed" gpurun_out/pytest_checked.log
FT="timeout 300 python tools/frame_time.py c4"
ARGS=""
for w in 1 2 4 8; do for sh in 1 0; do ARGS="$ARGS -- world=$w share=$sh reps=4"; done; done
$FT reps=4 $ARGS -- world=8 share=1 tail_pct=0 -- world=8 share=1 tail_pct=0 want_light=2 -- share=1 tail_pct=0 -- share=1 tail_pct=2 -- world=8 share=1 want_heavy_jobs=1 -- share=1 want_heavy_jobs=1 > gpurun_out/ab9_share_c4.log 2>&1; cat gpurun_out/ab9_share_c4.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 400 python tools/explore_tile.py 8 > gpurun_out/explore_tile_w8.log 2>&1; head -8 gpurun_out/explore_tile_w8.log
