#!/bin/bash
# AddressSanitizer + UndefinedBehaviorSanitizer over the library's HOST code, without a GPU: every .cu / .cc of csrc is
# compiled with -fsanitize=address,undefined (nvcc passes it to gcc; device code is unaffected and never runs), linked
# against the fake CUDA runtime of tests/fake_cuda/, and the host-logic, text-path, codec and flag tests run on it
# (300 rounds of random traffic with injected CUDA failures among them).  Usage: bash tools/host_asan.sh [out_dir]
set -e
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
OUT=${1:-/tmp/pgen_asan}
mkdir -p "$OUT"
cd "$ROOT/textsummarization-on-flink_b200/csrc"
SAN="-fsanitize=address -fsanitize=undefined -fno-omit-frame-pointer"
NV="-gencode arch=compute_100a,code=sm_100a -O1 -g -std=c++17 -Xcompiler -fPIC -DPGEN_LIVE_VT -DPGEN_LIVE_AF -DPGEN_LIVE_PE"
for s in $SAN; do NV="$NV -Xcompiler $s"; done
pids=()
for f in kernels engine capi noalloc vocab_tc attn_fused dec_cell; do /usr/local/cuda/bin/nvcc $NV -c $f.cu -o "$OUT/$f.o" & pids+=($!); done
for f in text tfexample; do g++ -std=c++17 -O1 -g -fPIC $SAN -c $f.cc -o "$OUT/$f.o" & pids+=($!); done
for p in "${pids[@]}"; do wait $p; done
cd "$ROOT"
gcc -c -fPIC -O1 -g -fsanitize=address -I/usr/local/cuda/include tests/fake_cuda/fake_cudart.c -o "$OUT/fake_cudart.o"
g++ -shared -Wl,-Bsymbolic $SAN -o "$OUT/libpgen_fake_asan.so" "$OUT"/*.o -lpthread
export PGEN_LIB="$OUT/libpgen_fake_asan.so" PGEN_FAKE_LIB="$OUT/libpgen_fake_asan.so" PGEN_FUZZ_ROUNDS=300
export LD_PRELOAD="$(gcc -print-file-name=libasan.so) $(gcc -print-file-name=libubsan.so)"
export ASAN_OPTIONS=detect_leaks=0:halt_on_error=1 UBSAN_OPTIONS=print_stacktrace=1
# (the two tests that expect a machine WITHOUT devices are left out: the fake runtime has two)
python -m pytest tests/test_host_logic_fake_cuda.py tests/test_reference_golden.py tests/test_tfexample_codec.py tests/test_capi_cpu.py \
  -q -s -m "not gpu" -p no:cacheprovider -k "not generator and not shim_ and not oracle and not no_cpu_fallback and not without_a_gpu" > "$OUT/run.log" 2>&1 || true
echo "runtime errors (UBSan): $(grep -c 'runtime error' "$OUT/run.log" || true)"
echo "AddressSanitizer reports: $(grep -c 'ERROR: AddressSanitizer' "$OUT/run.log" || true)"
tail -1 "$OUT/run.log"
