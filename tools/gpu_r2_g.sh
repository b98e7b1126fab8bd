#!/bin/bash
# round 2, seventh GPU pass (gpurun --gpus N): multi-GPU tests through the C ABI, bench lines at 1..N GPUs
# (frame_crc32 must agree), and on GPU 0 the A/B of the table staging modes and of want_heavy.
set -u
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L
FT="timeout 300 python tools/frame_time.py c4"
( $FT reps=4 -- world=8 reps=4 -- want_heavy=2 -- want_heavy=4 -- want_heavy=6 -- want_heavy=12 -- world=2 -- world=4
  for v in bulk1 sync; do $FT lib=build/variants/$v/librt_b200.so reps=4 -- world=8 reps=4; done ) > gpurun_out/ab7_c4.log 2>&1
echo "ab rc=$?"; cat gpurun_out/ab7_c4.log
RT_B200_TEST_LIB=build/variants/checked/librt_b200.so timeout 600 python -m pytest tests/test_gpu_sweep.py tests/test_gpu_parity.py -x -q -k "not true_size and not true_resolution" > gpurun_out/pytest_checked.log 2>&1; echo "pytest checked rc=$?"; tail -n 3 gpurun_out/pytest_checked.log
bash tools/gpu_r2_multi.sh $N
