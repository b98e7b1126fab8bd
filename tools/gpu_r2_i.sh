#!/bin/bash
# round 2, ninth GPU pass: grid over the heavy-batch slicing of tree sweeps (want_heavy) and shadow sweeps
# (want_heavy_jobs), with and without the super level, then the tail rule at 1 .. 1/8 of the frame.
set -u
mkdir -p gpurun_out
FT="timeout 600 python tools/frame_time.py c4"
ARGS=""
for t in 2 4 8 16; do for s in 1 2 4 8 16; do ARGS="$ARGS -- want_heavy=$t want_heavy_jobs=$s reps=4"; done; done
( $FT reps=4 $ARGS; $FT lib=build/variants/nosupers/librt_b200.so reps=4 $ARGS ) > gpurun_out/grid_heavy_c4.log 2>&1; cat gpurun_out/grid_heavy_c4.log
ARGS=""
for w in 1 2 4 8; do for m in 0 8 16 32 64; do ARGS="$ARGS -- world=$w tail_min=$m reps=4"; done; done
$FT reps=4 $ARGS > gpurun_out/grid_tail_c4.log 2>&1; cat gpurun_out/grid_tail_c4.log
