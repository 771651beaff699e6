"""Builds the library's own objects (csrc/build/*.o) against the fake CUDA runtime of this directory and returns a
`native` module bound to it plus the ctypes handle with the fake's control functions.  Test infrastructure only."""
import ctypes as C
import importlib.util
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CSRC = os.path.join(ROOT, "textsummarization-on-flink_b200", "csrc")


def build(out_dir):
    from textsummarization_on_flink_b200 import native as real_native
    objs = sorted(os.path.join(CSRC, "build", f) for f in os.listdir(os.path.join(CSRC, "build")) if f.endswith(".o"))
    assert len(objs) >= 9, "run __graft_entry__.build() first"
    fo, so = os.path.join(out_dir, "fake_cudart.o"), os.path.join(out_dir, "libpgen_fake.so")
    prebuilt = os.environ.get("PGEN_FAKE_LIB")       # e.g. an AddressSanitizer build of the same sources (tools/host_asan.sh)
    if prebuilt:
        so = prebuilt
    for cmd in () if prebuilt else (["gcc", "-c", "-fPIC", "-O1", "-I/usr/local/cuda/include", os.path.join(ROOT, "tests", "fake_cuda", "fake_cudart.c"), "-o", fo],
                ["g++", "-shared", "-Wl,-Bsymbolic", "-o", so] + objs + [fo, "-lpthread"]):
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    spec = importlib.util.spec_from_file_location("native_on_fake_cuda", real_native.__file__)
    nat = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(nat)
    nat.LIB_PATH = so
    L = nat.lib()
    L.fake_cuda_fail.argtypes = [C.c_char_p, C.c_int]
    for f in ("fake_cuda_launches", "fake_cuda_graph_launches", "fake_cuda_live_allocs", "fake_cuda_h2d_count", "fake_cuda_log_count"):
        getattr(L, f).restype = C.c_long
    L.fake_cuda_h2d_dst.restype, L.fake_cuda_h2d_dst.argtypes = C.c_void_p, [C.c_long]
    L.fake_cuda_h2d_bytes.restype, L.fake_cuda_h2d_bytes.argtypes = C.c_size_t, [C.c_long]
    L.fake_cuda_log_name.restype, L.fake_cuda_log_name.argtypes = C.c_char_p, [C.c_long]
    L.fake_cuda_log_dims.argtypes = [C.c_long, C.POINTER(C.c_uint * 5), C.POINTER(C.c_size_t)]
    L.fake_cuda_set_devices(2)
    return nat, L


def launch_log(L):
    """-> [(demangled kernel name without arguments, (gx, gy, gz), block, dynamic smem)]"""
    n = L.fake_cuda_log_count()
    names = [L.fake_cuda_log_name(i).decode() for i in range(n)]
    dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    out = []
    for i in range(n):
        d, sm = (C.c_uint * 5)(), C.c_size_t()
        L.fake_cuda_log_dims(i, C.byref(d), C.byref(sm))
        name = dem[i].replace("(anonymous namespace)::", "").replace("pgn::", "").replace("void ", "")
        name = name.split("(")[0]
        out.append((name, (d[0], d[1], d[2]), d[3], sm.value))
    return out
