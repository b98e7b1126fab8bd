#!/bin/bash
# round 2, tenth GPU pass: the share of a launch's batches that is cut finer at its end (tail_pct, tail_mult)
# at 1, 1/2, 1/4, 1/8 of the frame.
set -u
mkdir -p gpurun_out
FT="timeout 600 python tools/frame_time.py c4"
ARGS=""
for w in 1 2 4 8; do for p in 0 2 4 8 15; do for m in 2 4; do ARGS="$ARGS -- world=$w tail_pct=$p tail_mult=$m reps=4"; done; done; done
$FT reps=4 $ARGS > gpurun_out/grid_tail2_c4.log 2>&1; cat gpurun_out/grid_tail2_c4.log
ARGS=""
for w in 1 8; do for l in 1 2 3 4 6; do ARGS="$ARGS -- world=$w want_light=$l reps=4"; done; done
$FT reps=4 $ARGS > gpurun_out/grid_want_c4.log 2>&1; cat gpurun_out/grid_want_c4.log
