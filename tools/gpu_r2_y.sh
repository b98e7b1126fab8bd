#!/bin/bash
# round 2, GPU pass Y: ncu launch list of ONE 1/8 share of the c4 frame (serialised kernels: their sum against the share's frame time).
set -u
mkdir -p gpurun_out
timeout 200 python tools/frame_time.py c4 world=8 reps=3 > gpurun_out/share8_plain.log 2>&1; cat gpurun_out/share8_plain.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/share8_launches.csv \
    python tools/frame_time.py c4 world=8 reps=2 > gpurun_out/share8_ncu.log 2>&1; echo "ncu rc=$?"
