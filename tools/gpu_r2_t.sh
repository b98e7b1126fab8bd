#!/bin/bash
# round 2, GPU pass T: super pairs of a run visited nearest first (with the distance clip): tests, then near-first on / off.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sweep.py tests/test_gpu_forest.py tests/test_gpu_parity.py tests/test_gpu_grazing.py -x -q -k "not true_size" > gpurun_out/pytest_near.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_near.log | cut -c1-300
( timeout 300 python tools/frame_time.py c4 reps=4 -- world=2 reps=4 -- world=8 reps=4
  timeout 300 python tools/frame_time.py c4 lib=build/variants/nonear/librt_b200.so reps=4 -- world=8 reps=4
  timeout 200 python tools/frame_time.py c3 reps=4; timeout 200 python tools/frame_time.py c3 lib=build/variants/nonear/librt_b200.so reps=4 ) > gpurun_out/ab_near_first_c4.log 2>&1; cat gpurun_out/ab_near_first_c4.log
timeout 200 python tools/frame_time.py c4 reps=2 trace=1 2>&1 | grep "sweep [0-9]*:" | tail -25 | cut -c1-120
