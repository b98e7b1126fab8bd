#!/bin/bash
# round 2, third GPU pass: the GPU suite on the reworked sweep prologue / node order / filtered forest eval,
# A/B of the node order's cell size and the table staging, bench lines, then ncu (launch list + full capture).
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 16 gpurun_out/pytest_gpu.log
timeout 300 python -m pytest tests/test_gpu_grazing.py -q -s > gpurun_out/grazing.log 2>&1; grep "^case" gpurun_out/grazing.log
FT="timeout 200 python tools/frame_time.py c4"
( $FT -- node_bits=5 -- node_bits=6 -- node_bits=7 -- world=8 -- world=8 node_bits=5 -- world=8 node_bits=6 -- world=8 node_bits=7 -- shadow_anyhit=0 -- world=2 -- world=4
  $FT lib=build/variants/sync_table/librt_b200.so -- world=8 ) > gpurun_out/ab3_c4.log 2>&1
echo "ab rc=$?"; cat gpurun_out/ab3_c4.log
timeout 200 python tools/frame_time.py c3 -- node_bits=5 > gpurun_out/ab3_c3.log 2>&1; cat gpurun_out/ab3_c3.log
timeout 300 python bench.py > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench c4 rc=$?"; cut -c1-700 gpurun_out/bench_c4.json; tail -n 3 gpurun_out/bench_c4.err
timeout 120 python bench.py --workload c2f > gpurun_out/bench_c2f.json 2> gpurun_out/bench_c2f.err; echo "bench c2f rc=$?"; cat gpurun_out/bench_c2f.json; tail -n 3 gpurun_out/bench_c2f.err
bash tools/gpu_r2_ncu.sh
