#!/bin/bash
# Builds mobile_slam_b200/libvs_tracker.so for sm_100a (in-tree; the .so travels to the GPU box).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden -diag-suppress 550"
mkdir -p ../_obj
pids=()
for f in clahe pyramid lk shitomasi glue ransac tracker ingest no_alloc feature_bank; do
  if [ ! -f ../_obj/$f.o ] || [ $f.cu -nt ../_obj/$f.o ] || [ -n "$(find . -name '*.cuh' -newer ../_obj/$f.o -o -name '*.h' -newer ../_obj/$f.o)" ] || [ ../../include/vs_tracker.h -nt ../_obj/$f.o ]; then
    $NVCC $FLAGS -c $f.cu -o ../_obj/$f.o &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
$NVCC -shared -o ../libvs_tracker.so ../_obj/*.o -gencode arch=compute_100a,code=sm_100a -lz
echo "built $(realpath ../libvs_tracker.so)"
