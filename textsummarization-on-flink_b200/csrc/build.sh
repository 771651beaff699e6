#!/bin/bash
# Builds libpgen_cuda.so in-tree for sm_100a.  nvcc cross-compiles without a GPU.
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden"
# Kernel entry / exit counters (read by pgen_debug_where) are ON in the shipped library: a few fire-and-forget global
# reductions per CTA of the four kernels that wait on mbarriers.  DESIGN.md section 6 says why they stay in.
# PGEN_NO_LIVE=1 builds without them, PGEN_LIVE_ONLY=VT|AF|PE with one group only; PGEN_CYCLES=1 adds the clock64 phase
# counters of the attention kernel (tools/af_dbg.py).  Switching any of these needs `rm -rf build` first.
if [ -n "$PGEN_LIVE_ONLY" ]; then FLAGS="$FLAGS -DPGEN_LIVE_$PGEN_LIVE_ONLY"
elif [ -z "$PGEN_NO_LIVE" ]; then FLAGS="$FLAGS -DPGEN_LIVE_VT -DPGEN_LIVE_AF -DPGEN_LIVE_PE"; fi
[ -n "$PGEN_CYCLES" ] && FLAGS="$FLAGS -DPGEN_AF_CYCLES"
OUT=${PGEN_OUT:-../libpgen_cuda.so}
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
