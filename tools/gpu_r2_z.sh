#!/bin/bash
# round 2, last GPU pass: the c4 bench line and the ncu evidence of the committed code, and the quick GPU tests once more.
set -u
mkdir -p gpurun_out
timeout 300 python bench.py > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench c4 rc=$?"; cut -c1-200 gpurun_out/bench_c4.json
timeout 600 python -m pytest tests/test_gpu_sweep.py tests/test_gpu_forest.py tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q -k "not true_size and not true_resolution" > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"; tail -n 2 gpurun_out/pytest_quick.log
bash tools/gpu_r2_ncu.sh > gpurun_out/ncu_pass.log 2>&1; grep "rc=" gpurun_out/ncu_pass.log
