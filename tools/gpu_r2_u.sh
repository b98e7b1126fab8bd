#!/bin/bash
# round 2, GPU pass U: slice-major items that start from the keys earlier slices left: tests, then on / off at 1 .. 1/8 of the frame.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sweep.py tests/test_gpu_forest.py tests/test_gpu_parity.py tests/test_gpu_grazing.py -x -q -k "not true_size" > gpurun_out/pytest_slicemajor.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_slicemajor.log | cut -c1-300
( timeout 300 python tools/frame_time.py c4 reps=4 -- world=2 reps=4 -- world=4 reps=4 -- world=8 reps=4 -- world=8 rank=7 reps=4
  timeout 300 python tools/frame_time.py c4 lib=build/variants/batchmajor/librt_b200.so reps=4 -- world=2 reps=4 -- world=4 reps=4 -- world=8 reps=4 -- world=8 rank=7 reps=4 ) > gpurun_out/ab_slice_major_c4.log 2>&1; cat gpurun_out/ab_slice_major_c4.log
