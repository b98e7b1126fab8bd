#!/bin/bash
# round 2, fifth GPU pass: scheduling-knob sweep at 1 and 1/8 of the frame, per-sweep trace of both,
# compile-time variants of the sweep (quad pairs two at a time, brute-loop unroll, flush threshold, table staging).
set -u
mkdir -p gpurun_out
FT="timeout 300 python tools/frame_time.py c4"
( $FT reps=3 -- item_max=8 reps=3 -- item_max=12 reps=3 -- item_max=24 reps=3 -- item_max=32 reps=3 -- want_light=2 reps=3 -- want_light=8 reps=3 -- tail_pct=0 reps=3 -- tail_pct=30 reps=3 -- tail_mult=8 reps=3 -- want_heavy=24 reps=3 -- want_heavy=96 reps=3 -- heavy_allow=0 reps=3 -- merge_lights=0 reps=3
  $FT world=8 reps=3 -- world=8 item_max=8 reps=3 -- world=8 item_max=24 reps=3 -- world=8 want_light=2 reps=3 -- world=8 want_light=8 reps=3 -- world=8 tail_pct=0 reps=3 -- world=8 tail_pct=30 reps=3 -- world=8 tail_mult=8 reps=3 -- world=8 want_heavy=96 reps=3 -- world=8 node_bits=7 reps=3 -- world=8 rays_per_thread=2 reps=3 -- world=8 merge_lights=0 reps=3 ) > gpurun_out/knobs_c4.log 2>&1
echo "knobs rc=$?"; cat gpurun_out/knobs_c4.log
for v in quad2 brute1 brute4 flush64 sync_table; do $FT lib=build/variants/$v/librt_b200.so reps=3 -- world=8 reps=3; done > gpurun_out/ab5_c4.log 2>&1; cat gpurun_out/ab5_c4.log
$FT trace=1 reps=1 > gpurun_out/trace_w1.out 2> gpurun_out/trace_w1.log; $FT world=8 trace=1 reps=1 > gpurun_out/trace_w8.out 2> gpurun_out/trace_w8.log
grep "sweep " gpurun_out/trace_w1.log | head -20; grep "sweep " gpurun_out/trace_w8.log | head -20
ls -la gpurun_out
