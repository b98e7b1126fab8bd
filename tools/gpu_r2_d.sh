#!/bin/bash
# round 2, fourth GPU pass: the smaller sweep code (rolled cold loops) against the previous build, quick
# correctness on it, and the launch list of one rank's 1/8 share of the frame.
set -u
mkdir -p gpurun_out
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/smoke.log
timeout 600 python -m pytest tests/test_gpu_sweep.py tests/test_gpu_grazing.py tests/test_gpu_forest.py -x -q > gpurun_out/pytest_quick.log 2>&1; echo "pytest quick rc=$?"; tail -n 4 gpurun_out/pytest_quick.log
FT="timeout 200 python tools/frame_time.py c4"
( $FT -- world=8 -- world=4 -- world=8 node_bits=7
  $FT lib=build/variants/prev/librt_b200.so -- world=8 -- world=4 ) > gpurun_out/ab4_c4.log 2>&1
echo "ab rc=$?"; cat gpurun_out/ab4_c4.log
( timeout 100 python tools/frame_time.py c3; timeout 100 python tools/frame_time.py c3 lib=build/variants/prev/librt_b200.so;
  timeout 100 python tools/frame_time.py c2; timeout 100 python tools/frame_time.py c2 lib=build/variants/prev/librt_b200.so ) > gpurun_out/ab4_c3.log 2>&1; cat gpurun_out/ab4_c3.log
CMD="python tools/frame_time.py c4 world=8 reps=3"
timeout 120 $CMD > gpurun_out/share8_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/share8_launches.csv $CMD > gpurun_out/share8_ncu.log 2>&1; echo "ncu share rc=$?"
ls -la gpurun_out
