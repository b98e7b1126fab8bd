#!/bin/bash
# round 2, GPU pass N: per-sweep trace of c4 at the whole frame and at a 1/8 share (serial sweeps, one span each),
# then the ncu launch list and the full capture of the sweep at HEAD.
set -u
mkdir -p gpurun_out
timeout 200 python tools/frame_time.py c4 reps=2 trace=1 > gpurun_out/trace_c4_w1.log 2>&1
timeout 200 python tools/frame_time.py c4 reps=2 trace=1 world=8 > gpurun_out/trace_c4_w8.log 2>&1
grep -c sweep gpurun_out/trace_c4_w1.log gpurun_out/trace_c4_w8.log
bash tools/gpu_r2_ncu.sh
