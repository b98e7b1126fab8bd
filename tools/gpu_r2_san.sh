#!/bin/bash
# One gpurun call with ONE compute-sanitizer tool (memcheck | racecheck | initcheck | synccheck) over
# smoke() and the tests of the sweep machinery that drive every path of the kernel on small frames.
#   bash tools/gpu_r2_san.sh memcheck
set -u
TOOL=${1:-memcheck}
mkdir -p gpurun_out
timeout 900 compute-sanitizer --tool $TOOL --error-exitcode 9 --log-file gpurun_out/sanitizer_${TOOL}_smoke.log \
    python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/sanitizer_${TOOL}_smoke.out 2>&1; echo "$TOOL smoke rc=$?"
tail -n 3 gpurun_out/sanitizer_${TOOL}_smoke.out; tail -n 4 gpurun_out/sanitizer_${TOOL}_smoke.log
timeout 1500 compute-sanitizer --tool $TOOL --error-exitcode 9 --log-file gpurun_out/sanitizer_${TOOL}_sweep.log \
    python -m pytest tests/test_gpu_sweep.py -x -q -k "chunked or queue or merged or work_item or bands" \
    > gpurun_out/sanitizer_${TOOL}_sweep.out 2>&1; echo "$TOOL sweep tests rc=$?"
tail -n 4 gpurun_out/sanitizer_${TOOL}_sweep.out; tail -n 4 gpurun_out/sanitizer_${TOOL}_sweep.log
ls -la gpurun_out
