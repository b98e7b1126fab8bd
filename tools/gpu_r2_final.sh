#!/bin/bash
# round 2, final GPU pass: the whole GPU suite, smoke, frame shares, every bench line, the reference arm, then ncu
# (launch list + full capture of the sweep) of the same bench command.
set -u
bash tools/gpu_r2_m.sh
bash tools/gpu_r2_ncu.sh
