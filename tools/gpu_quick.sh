#!/bin/bash
# A short gpurun pass (a few minutes of box time): the c4 bench line, the GPU tests, smoke, and the
# ncu launch list of the same bench command.  Each step has its own timeout; outputs in gpurun_out/.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap --format=csv -lms 500 > gpurun_out/clocks.csv &
SMI=$!
timeout 240 python bench.py > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; rc=$?; echo "bench c4 rc=$rc"; cat gpurun_out/bench_c4.json
kill $SMI
timeout 330 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 4 gpurun_out/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/smoke.log
if [ $rc -eq 0 ]; then
  timeout 150 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
      python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
fi
if [ $rc -eq 0 ] && [ "${FULL:-0}" = "1" ]; then
  # one full capture of the dominant kernel: the level-1 tree sweep (general filter) and the level-0 shadow sweep (all lights, one launch)
  timeout 240 ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -s 1 -c 2 -f -o gpurun_out/sweep_full \
      python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
fi
timeout 60 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2>&1; echo "bench ref rc=$?"; cat gpurun_out/bench_ref.json
ls -la gpurun_out
