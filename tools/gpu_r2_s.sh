#!/bin/bash
# round 2, GPU pass S: distance clip of the nearest-hit sweeps (spheres of the hierarchy wholly beyond a ray's best hit so far are
# not descended into): sweep / forest / parity / grazing tests, then the frame times.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sweep.py tests/test_gpu_forest.py tests/test_gpu_parity.py tests/test_gpu_grazing.py -x -q -k "not true_size" > gpurun_out/pytest_clip.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_clip.log | cut -c1-300
timeout 300 python tools/frame_time.py c4 reps=4 -- world=8 reps=4 > gpurun_out/ab_clip_c4.log 2>&1; cat gpurun_out/ab_clip_c4.log
timeout 200 python tools/frame_time.py c3 reps=4 2>&1 | tail -1
timeout 200 python tools/frame_time.py c4 reps=2 trace=1 2>&1 | grep "sweep [0-9]*:" | tail -25
