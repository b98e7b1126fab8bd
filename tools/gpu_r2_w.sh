#!/bin/bash
# round 2, GPU pass W: primary rays in tiles of 16 x 8 pixels: tests (parity incl. true resolution rows, bands, ragged frames, forest), on / off.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_sweep.py tests/test_gpu_forest.py tests/test_gpu_parity.py tests/test_cli.py -x -q > gpurun_out/pytest_tiles.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_tiles.log | cut -c1-300
ARGS=""
for w in 1 8; do for v in 1 0; do ARGS="$ARGS -- world=$w primary_tiles=$v reps=4"; done; done
timeout 400 python tools/frame_time.py c4 reps=2 $ARGS > gpurun_out/ab_primary_tiles_c4.log 2>&1; cat gpurun_out/ab_primary_tiles_c4.log
timeout 200 python tools/frame_time.py c3 reps=4 -- primary_tiles=0 reps=4 2>&1 | tail -2
timeout 200 python tools/frame_time.py c2 reps=4 -- primary_tiles=0 reps=4 2>&1 | tail -2
timeout 200 python tools/frame_time.py c4 reps=2 trace=1 2>&1 | grep "sweep 0:"
