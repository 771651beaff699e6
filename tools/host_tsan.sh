#!/bin/bash
# ThreadSanitizer over the library's host code: handles on several devices driven from several threads of one process
# (tests/fake_cuda/threads_driver.cc), built with -fsanitize=thread against the fake CUDA runtime.  No GPU needed.
# Usage: bash tools/host_tsan.sh [out_dir]
set -e
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
OUT=${1:-/tmp/pgen_tsan}
mkdir -p "$OUT"
cd "$ROOT/textsummarization-on-flink_b200/csrc"
NV="-gencode arch=compute_100a,code=sm_100a -O1 -g -std=c++17 -Xcompiler -fPIC -Xcompiler -fsanitize=thread -DPGEN_LIVE_VT -DPGEN_LIVE_AF -DPGEN_LIVE_PE"
pids=()
for f in kernels engine capi noalloc vocab_tc attn_fused dec_cell; do /usr/local/cuda/bin/nvcc $NV -c $f.cu -o "$OUT/$f.o" & pids+=($!); done
for f in text tfexample; do g++ -std=c++17 -O1 -g -fPIC -fsanitize=thread -c $f.cc -o "$OUT/$f.o" & pids+=($!); done
for p in "${pids[@]}"; do wait $p; done
cd "$ROOT"
gcc -c -fPIC -O1 -g -fsanitize=thread -I/usr/local/cuda/include tests/fake_cuda/fake_cudart.c -o "$OUT/fake_cudart.o"
g++ -std=c++17 -O1 -g -fsanitize=thread -Iinclude tests/fake_cuda/threads_driver.cc "$OUT"/*.o -lpthread -o "$OUT/threads_driver"
MODEL=$(python - <<'PY'
import os, sys, tempfile
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import pgen_pkg; pgen_pkg.load()
import numpy as np
from textsummarization_on_flink_b200 import synth, weights_io
from conftest import small_flags
V, H, E = 300, 32, 16
root = tempfile.mkdtemp(prefix="pgen_tsan_model_")
train = os.path.join(root, "log", "small", "train"); os.makedirs(train)
weights_io.save_weights(os.path.join(train, weights_io.DEFAULT_NAME), synth.make_weights("W1", vsize=V, hidden=H, emb=E, seed=7))
synth.write_vocab_file(os.path.join(root, "vocab"), V)
print(small_flags({"vocab_path": os.path.join(root, "vocab"), "log_root": os.path.join(root, "log"), "exp_name": "small"}, dict(V=V, H=H, E=E),
                  extra="--articles_per_batch=4 --slots=2 --flush_ms=-1"))
PY
)
TSAN_OPTIONS=halt_on_error=0 "$OUT/threads_driver" "$MODEL" 8 > "$OUT/run.log" 2>&1 || true
echo "ThreadSanitizer reports: $(grep -c 'WARNING: ThreadSanitizer' "$OUT/run.log" || true)"
tail -1 "$OUT/run.log"
