#!/bin/bash
# round 2, GPU pass Q: one full ncu capture of the wavefront kernels between the sweeps (spawn, shade, node order) on c4.
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain_for_ncu2.json 2> gpurun_out/plain_for_ncu2.err &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"spawn_kernel|shade_kernel|node_bin_kernel|node_scatter_kernel" -s 0 -c 8 -f -o gpurun_out/wavefront_full \
    $CMD > gpurun_out/ncu_full2.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out | tail -5
