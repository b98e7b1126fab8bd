#!/bin/bash
# round 2, second GPU pass: A/B of the sweep variants and of the node order's cell size, then (ncu) the
# launch list and one full capture of the dominant kernel at HEAD.
set -u
mkdir -p gpurun_out
FT="timeout 200 python tools/frame_time.py c4"
( $FT -- shadow_anyhit=0 -- node_bits=6 -- node_bits=7 -- node_bits=4 -- world=8 -- world=8 shadow_anyhit=0 -- world=8 node_bits=6 -- world=8 node_bits=7
  $FT lib=build/variants/coef_once/librt_b200.so -- shadow_anyhit=0 -- world=8
  $FT lib=build/variants/sync_table/librt_b200.so -- world=8 ) > gpurun_out/ab2_c4.log 2>&1
echo "ab rc=$?"; cat gpurun_out/ab2_c4.log
timeout 200 python tools/frame_time.py c3 -- node_bits=6 -- node_bits=7 > gpurun_out/ab2_c3.log 2>&1; cat gpurun_out/ab2_c3.log
bash tools/gpu_r2_ncu.sh
