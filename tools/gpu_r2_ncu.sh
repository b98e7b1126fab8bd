#!/bin/bash
# One gpurun call with ncu only: the launch list of the bench command and one full capture of the
# dominant kernel (sweep launches 1 and 2 of the run: the level-1 tree sweep with the general filter
# and the level-0 shadow sweep of all lights).  The same command runs plain first.
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain_for_ncu.json 2> gpurun_out/plain_for_ncu.err &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
    $CMD > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -s 1 -c 2 -f -o gpurun_out/sweep_full \
    $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
cat gpurun_out/plain_for_ncu.json
ls -la gpurun_out
