#!/bin/bash
# Builds an A/B variant of libptb200.so with extra -D flags for kernels_scene.cu:  experiments/build_variant.sh NAME -DFOO=1 ...
# -> experiments/variants/libptb200_NAME.so (select with PTB200_LIB=...)
set -e
cd "$(dirname "$0")/../pathtracing_b200/csrc"
name=$1; shift
mkdir -p ../../experiments/variants
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr -fmad=false "$@" \
     -Xptxas -v -c -o /tmp/kernels_scene_$name.o kernels_scene.cu 2> /tmp/kernels_scene_$name.log
grep -A2 "bounce_kernelILb0ELb0" /tmp/kernels_scene_$name.log | grep -E "registers|spill" || true
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../experiments/variants/libptb200_$name.so api.o kernels_csmain.o kernels_image.o selftest.o scene.o scene_build.o /tmp/kernels_scene_$name.o -cudart shared
