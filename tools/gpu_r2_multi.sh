#!/bin/bash
# One gpurun --gpus N call: the multi-GPU tests (rt_multi through the C ABI, rt_cli --gpus) and the
# strong-scaling bench lines at 1 .. N GPUs (frame_crc32 must be the same in every line).
set -u
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_cli.py -x -q -k "multi_gpu or gpus_flag or frame_rows or frame_signal" \
    > gpurun_out/pytest_multi_n$N.log 2>&1; echo "pytest multi rc=$?"; tail -n 5 gpurun_out/pytest_multi_n$N.log
for n in 1 2 4 8; do
  if [ $n -gt $N ]; then break; fi
  if [ $n -eq 1 ]; then
    timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_c4_n$n.json 2> gpurun_out/bench_c4_n$n.err
  else
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
        bench.py --gpus $n > gpurun_out/bench_c4_n$n.json 2> gpurun_out/bench_c4_n$n.err
  fi
  echo "bench n=$n rc=$?"; cat gpurun_out/bench_c4_n$n.json | cut -c1-400; tail -n 3 gpurun_out/bench_c4_n$n.err
done
grep -o '"frame_crc32": [0-9]*' gpurun_out/bench_c4_n*.json
ls -la gpurun_out
