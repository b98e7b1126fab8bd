#!/bin/bash
# round 2, GPU pass M: the no-allocation entry point alone and after the suite, the whole GPU suite at HEAD
# (sweep without work sharing), smoke, frame times, the bench lines of every workload and of the reference arm.
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/pytest_no_alloc_alone.log 2>&1; echo "no_alloc alone rc=$?"; tail -n 3 gpurun_out/pytest_no_alloc_alone.log | cut -c1-300
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 14 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 300 python tools/frame_time.py c4 reps=4 -- world=2 -- world=4 -- world=8 > gpurun_out/frame_c4.log 2>&1; cat gpurun_out/frame_c4.log
timeout 300 python bench.py > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench c4 rc=$?"; cut -c1-900 gpurun_out/bench_c4.json; tail -n 3 gpurun_out/bench_c4.err
for w in c2 c3 c5 c2f; do
  timeout 400 python bench.py --workload $w > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"; cut -c1-500 gpurun_out/bench_$w.json; tail -n 2 gpurun_out/bench_$w.err
done
timeout 400 python bench.py --impl reference > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "bench ref rc=$?"; cut -c1-500 gpurun_out/bench_reference.json
