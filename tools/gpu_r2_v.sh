#!/bin/bash
# round 2, GPU pass V: a level's shadow sweep is not joined before the next level starts (level_pipeline): tests, on / off.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sweep.py tests/test_gpu_forest.py tests/test_gpu_parity.py tests/test_cli.py tests/test_zz_no_alloc.py -x -q -k "not true_size" > gpurun_out/pytest_pipeline.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_pipeline.log | cut -c1-300
ARGS=""
for w in 1 2 4 8; do for v in 1 0; do ARGS="$ARGS -- world=$w level_pipeline=$v reps=4"; done; done
timeout 400 python tools/frame_time.py c4 reps=2 $ARGS > gpurun_out/ab_level_pipeline_c4.log 2>&1; cat gpurun_out/ab_level_pipeline_c4.log
timeout 200 python tools/frame_time.py c3 reps=4 -- level_pipeline=0 reps=4 2>&1 | tail -2
timeout 200 python tools/frame_time.py c5 reps=2 -- level_pipeline=0 reps=2 2>&1 | tail -2
