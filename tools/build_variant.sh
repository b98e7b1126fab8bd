#!/bin/bash
# Build a variant of librt_b200.so with extra nvcc defines, for A/B runs on the GPU box:
#   tools/build_variant.sh NAME "-DRT_SWEEP_COEF_ONCE=1"
# -> build/variants/NAME/librt_b200.so (build/ is git-ignored but travels with gpurun);
#    python tools/frame_time.py c4 lib=build/variants/NAME/librt_b200.so
set -eu
NAME=$1; DEFS=${2:-}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
V=$ROOT/build/variants
mkdir -p $V/$NAME
rm -rf $V/$NAME/csrc
mkdir -p $V/include && cp $ROOT/include/*.h $V/include/
mkdir -p $V/$NAME/csrc
(cd $ROOT/rust_tracer_b200/csrc && cp -r *.cu *.cuh *.h Makefile host $V/$NAME/csrc/)
rm -f $V/$NAME/csrc/*.o $V/$NAME/csrc/host/*.o
make -C $V/$NAME/csrc NVCCFLAGS="-O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-ffp-contract=off -Xptxas -v $DEFS" ../librt_b200.so > $V/$NAME/build.log 2>&1 || (tail -20 $V/$NAME/build.log; exit 1)
rm -f $V/$NAME/csrc/*.o $V/$NAME/csrc/host/*.o
ls -la $V/$NAME/librt_b200.so
