#!/bin/bash
# One gpurun call: GPU tests, smoke, the bench lines, then (each only after its own command
# exited 0 without ncu) the ncu launch list of the same bench command and one full capture of
# the dominant kernel (launches 1 and 2 of the sweep: the level-1 tree sweep and the level-0 shadow sweep of all lights).
# Outputs land in gpurun_out/.  tools/gpu_quick.sh is the short version (c4 only; FULL=1 adds the capture).
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/smoke.log
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap --format=csv -lms 500 > gpurun_out/clocks.csv &
SMI=$!
python bench.py > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; rc=$?; echo "bench c4 rc=$rc"; cat gpurun_out/bench_c4.json
python bench.py --workload c3 --no-cpu-baseline > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench c3 rc=$?"; cat gpurun_out/bench_c3.json
python bench.py --workload c2 --no-cpu-baseline > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 rc=$?"; cat gpurun_out/bench_c2.json
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2>&1; echo "bench ref rc=$?"; cat gpurun_out/bench_ref.json
kill $SMI
if [ $rc -eq 0 ]; then
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/plain_for_ncu.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
      python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:sweep_kernel -s 1 -c 2 -f -o gpurun_out/sweep_full \
      python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
fi
ls -la gpurun_out
