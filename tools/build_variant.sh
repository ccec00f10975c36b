#!/bin/bash
# Build libicx.so with extra nvcc flags (profiling variants), e.g. tools/build_variant.sh -DICX_PROFILE_PAIR
# Restore the product build with: python __graft_entry__.py (after touching a source) or tools/build_variant.sh
set -e
cd "$(dirname "$0")/.."
mkdir -p build
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true -ftz=false -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -Wno-deprecated-gpu-targets"
pids=()
for s in icx_abi icx_kernels icx_resolve icx_convert icx_timeline icx_clone icx_behaviors icx_nccl icx_mesh; do
  nvcc $FLAGS "$@" -I include -c intercession_b200/csrc/$s.cu -o build/$s.cu.o & pids+=($!)
done
for p in "${pids[@]}"; do wait $p; done
nvcc -shared -o intercession_b200/libicx.so build/icx_abi.cu.o build/icx_kernels.cu.o build/icx_resolve.cu.o build/icx_convert.cu.o build/icx_timeline.cu.o build/icx_clone.cu.o build/icx_behaviors.cu.o build/icx_nccl.cu.o build/icx_mesh.cu.o -lcudart -ldl
echo built "$@"
