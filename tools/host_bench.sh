#!/bin/bash
# Host cost of the operator's row path per device batch, without a GPU: the library's objects on the fake CUDA runtime
# (tests/fake_cuda/), driven by tests/fake_cuda/host_bench.cc.  Usage: bash tools/host_bench.sh [batches]
set -e
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
OUT=/tmp/pgen_host_bench
mkdir -p "$OUT"
cd "$ROOT"
gcc -c -fPIC -O2 -I/usr/local/cuda/include tests/fake_cuda/fake_cudart.c -o "$OUT/fake_cudart.o"
g++ -std=c++17 -O2 -Iinclude tests/fake_cuda/host_bench.cc textsummarization-on-flink_b200/csrc/build/*.o "$OUT/fake_cudart.o" -lpthread -o "$OUT/host_bench"
FLAGS=$(python - <<'PY'
import os, sys
sys.path.insert(0, os.getcwd())
import pgen_pkg; pgen_pkg.load()
from textsummarization_on_flink_b200 import synth
paths = synth.make_model_dir(os.environ.get("PGEN_BENCH_MODEL", "/tmp/pgen_bench_model"), "W1")
open("/tmp/pgen_host_bench/articles.txt", "wb").write(b"\n".join(synth.make_articles(1024, seed=1234)) + b"\n")
print(synth.hyperparams_string(paths, extra="--articles_per_batch=64 --flush_ms=-1"))
PY
)
"$OUT/host_bench" "$FLAGS" "$OUT/articles.txt" "${1:-40}"
