#!/bin/bash
# round 2, GPU pass P: CTAs of the all-lights shadow launch go on with another light's items when their own run out
# (job_rotate): sweep / parity tests, then on / off at 1 .. 1/8 of the frame.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sweep.py tests/test_gpu_forest.py tests/test_gpu_parity.py -x -q -k "not true_size" > gpurun_out/pytest_rotate.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_rotate.log | cut -c1-300
ARGS=""
for w in 1 2 4 8; do for v in 1 0; do ARGS="$ARGS -- world=$w job_rotate=$v reps=4"; done; done
timeout 400 python tools/frame_time.py c4 reps=4 $ARGS > gpurun_out/ab_job_rotate_c4.log 2>&1; cat gpurun_out/ab_job_rotate_c4.log
timeout 200 python tools/frame_time.py c3 reps=4 -- job_rotate=0 reps=4 > gpurun_out/ab_job_rotate_c3.log 2>&1; cat gpurun_out/ab_job_rotate_c3.log
timeout 200 python tools/frame_time.py c4 reps=2 trace=2 world=8 2>&1 | grep "sweep [0-9]\|frame" | tail -8
