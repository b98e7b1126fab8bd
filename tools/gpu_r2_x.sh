#!/bin/bash
# round 2, GPU pass X: the scheduling knobs again after the clip / slice-major / pipelined levels (whole frame and a 1/8 share).
set -u
mkdir -p gpurun_out
ARGS=""
for k in "want_light=2" "want_light=8" "tail_pct=0" "tail_pct=16" "tail_pct=16 tail_mult=4" "want_heavy_jobs=2" "want_heavy_jobs=8" "want_heavy=2" "want_heavy=8" "item_max=8" "item_max=24" "node_bits=6" "node_bits=7"; do
  ARGS="$ARGS -- $k reps=3 -- $k world=8 reps=3"
done
timeout 800 python tools/frame_time.py c4 reps=3 -- world=8 reps=3 $ARGS > gpurun_out/knob_sweep2_c4.log 2>&1; cat gpurun_out/knob_sweep2_c4.log
