#!/bin/bash
# round 2, first GPU pass: the reworked sweep (any-hit shadow rays, bulk-copied table) -- tests of the
# sweep machinery first, each step under its own timeout, then an A/B of the frame time, the whole GPU
# suite and the bench line.
set -u
mkdir -p gpurun_out
nvidia-smi -L
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/smoke.log
timeout 600 python -m pytest tests/test_gpu_sweep.py -x -q > gpurun_out/pytest_sweep.log 2>&1; echo "pytest sweep rc=$?"; tail -n 5 gpurun_out/pytest_sweep.log
timeout 300 python tools/frame_time.py c4 -- shadow_anyhit=0 -- world=8 rank=0 -- world=8 rank=0 shadow_anyhit=0 > gpurun_out/ab_c4.log 2>&1; echo "ab rc=$?"; cat gpurun_out/ab_c4.log
timeout 1500 python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 30 gpurun_out/pytest_gpu.log
timeout 300 python bench.py > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; rc=$?; echo "bench c4 rc=$rc"; cat gpurun_out/bench_c4.json; tail -n 5 gpurun_out/bench_c4.err
timeout 120 python bench.py --workload c2f > gpurun_out/bench_c2f.json 2> gpurun_out/bench_c2f.err; echo "bench c2f rc=$?"; cat gpurun_out/bench_c2f.json; tail -n 5 gpurun_out/bench_c2f.err
ls -la gpurun_out
