#!/bin/bash
# Builds libpgen_cuda.so in-tree for sm_100a.  nvcc cross-compiles without a GPU.
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden"
OUT=../libpgen_cuda.so
CU="kernels engine capi noalloc vocab_tc attn_fused dec_cell"
CC="text tfexample"
mkdir -p build
pids=()
for f in $CU; do
  o=build/$f.o
  if [ ! -f $o ] || [ $f.cu -nt $o ] || [ -n "$(find . ../../include -maxdepth 1 -name '*.h' -newer $o)" ]; then
    rm -f $o                                     # a failed compile must not leave a stale object to link
    $NVCC $FLAGS -c $f.cu -o $o &
    pids+=($!)
  fi
done
for f in $CC; do
  o=build/$f.o
  if [ ! -f $o ] || [ $f.cc -nt $o ] || [ $f.h -nt $o ]; then
    rm -f $o
    g++ -std=c++17 -O2 -fPIC -fvisibility=hidden -c $f.cc -o $o &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done          # `wait` without arguments would swallow a failed job's status
objs=""
for f in $CU $CC; do objs="$objs build/$f.o"; done
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o $OUT $objs -cudart static
echo built $OUT
