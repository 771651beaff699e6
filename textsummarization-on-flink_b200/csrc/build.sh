#!/bin/bash
# Builds libpgen_cuda.so in-tree for sm_100a.  nvcc cross-compiles without a GPU.
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden"
OUT=../libpgen_cuda.so
mkdir -p build
for f in kernels.cu engine.cu capi.cu noalloc.cu vocab_tc.cu; do
  if [ ! -f build/${f%.cu}.o ] || [ $f -nt build/${f%.cu}.o ] || [ -n "$(find . ../../include -maxdepth 1 -name '*.h' -newer build/${f%.cu}.o)" ]; then
    $NVCC $FLAGS -c $f -o build/${f%.cu}.o &
  fi
done
for f in text tfexample; do
  if [ ! -f build/$f.o ] || [ $f.cc -nt build/$f.o ] || [ $f.h -nt build/$f.o ]; then
    g++ -std=c++17 -O2 -fPIC -fvisibility=hidden -c $f.cc -o build/$f.o &
  fi
done
wait
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o $OUT build/kernels.o build/engine.o build/capi.o build/noalloc.o build/vocab_tc.o build/text.o build/tfexample.o -cudart static
echo built $OUT
