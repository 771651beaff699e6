"""The JNI shim of INTEGRATION.md section 3 (textsummarization-on-flink_b200/jni/pgen_jni.c): no JDK exists in this
image, so it is compiled against a STAND-IN jni.h (tests/jni_stub/jni.h: the JNI specification's signatures of the
members the shim uses, nothing else) with -Wall -Werror, linked with libpgen_cuda.so, and driven from Python with a
mock JNIEnv (tests/jni_stub/mock_jvm.c).  What runs here, without a GPU: deviceCount, destroy(0), and create's failure
path -- the status code becomes a pending java.lang.RuntimeException carrying pgen_last_error's message, like
TFModel.java:78-80 wraps every failure.  The submit / poll path needs a handle, i.e. a GPU: its marshalling runs against
a test double of the library (tests/jni_stub/fake_pgen.c); through the real library it has not been run."""
import ctypes as C
import os
import re
import subprocess

import pytest

from textsummarization_on_flink_b200 import native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "textsummarization-on-flink_b200", "jni", "pgen_jni.c")
STUB = os.path.join(ROOT, "tests", "jni_stub")
PKG = "Java_org_apache_flink_table_ml_lib_tensorflow_PgenNative_"


def test_integration_md_shows_the_shim_that_is_compiled():
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    block = re.search(r"## 3\. The JNI shim[^\n]*\n\n```c\n(.*?)```", doc, re.S).group(1)
    src = open(SHIM).read()
    assert src.endswith(block), "INTEGRATION.md section 3 and jni/pgen_jni.c differ"


@pytest.fixture(scope="module")
def shim(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("jni") / "libpgen_jni_test.so")
    libdir = os.path.dirname(native.LIB_PATH)
    cmd = ["gcc", "-shared", "-fPIC", "-std=gnu99", "-Wall", "-Wextra", "-Werror", "-Wno-unused-parameter", "-I", STUB,
           "-I", os.path.join(ROOT, "include"), SHIM, os.path.join(STUB, "mock_jvm.c"), "-L", libdir, "-lpgen_cuda",
           "-Wl,-rpath," + libdir, "-o", out]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    L = C.CDLL(out)
    L.mock_env.restype = C.c_void_p
    for f in ("mock_new_string", "mock_new_bytes", "mock_element"):
        getattr(L, f).restype = C.c_void_p
    L.mock_new_string.argtypes = [C.c_char_p]
    L.mock_new_bytes.argtypes = [C.c_char_p, C.c_int]
    L.mock_element.argtypes = [C.c_void_p, C.c_int]
    for f in ("mock_exception_class", "mock_exception_message"):
        getattr(L, f).restype = C.c_char_p
    L.mock_data.restype = C.c_void_p
    L.mock_data.argtypes = [C.c_void_p]
    L.mock_length.argtypes = [C.c_void_p]
    L.mock_borrowed.argtypes = [C.c_void_p]
    create = getattr(L, PKG + "create")
    create.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]
    create.restype = C.c_int64
    getattr(L, PKG + "deviceCount").argtypes = [C.c_void_p, C.c_void_p]
    getattr(L, PKG + "deviceCount").restype = C.c_int32
    getattr(L, PKG + "destroy").argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
    getattr(L, PKG + "destroy").restype = None
    return L


def test_shim_exports_the_five_native_methods(shim):
    for m in ("create", "submit", "poll", "destroy", "deviceCount"):       # PgenNative's natives, INTEGRATION.md section 2
        assert hasattr(shim, PKG + m), m


def test_device_count_and_destroy_of_nothing(shim):
    env = shim.mock_env()
    assert getattr(shim, PKG + "deviceCount")(env, None) == native.device_count()
    getattr(shim, PKG + "destroy")(env, None, 0)                           # pgen_destroy(NULL) is a no-op
    assert shim.mock_exception_class() is None


def test_create_failure_becomes_a_runtime_exception(shim, tmp_path):
    env = shim.mock_env()
    cases = [("p --mode=train", "mode=train"), ("p --mode=banana", "must be one of train/eval/decode"),
             ("p --mode=decode --vocab_path=%s" % (tmp_path / "missing"), "vocab")]
    for flags, needle in cases:
        shim.mock_clear_exception()
        js = shim.mock_new_string(flags.encode())
        h = getattr(shim, PKG + "create")(env, None, js, 0)
        assert h == 0
        assert shim.mock_exception_class() == b"java/lang/RuntimeException"
        assert needle in shim.mock_exception_message().decode(), shim.mock_exception_message()
        assert shim.mock_borrowed(js) == 0                                 # GetStringUTFChars was released
    shim.mock_clear_exception()


def test_mock_objects_behave(shim):
    b = shim.mock_new_bytes(b"a\x00bc", 4)
    assert shim.mock_length(b) == 4 and C.string_at(shim.mock_data(b), 4) == b"a\x00bc"


def test_submit_and_poll_marshalling_against_a_test_double(tmp_path):
    """The shim's byte[] -> C and rows -> byte[][][] marshalling, on the CPU: the library is replaced by a TEST DOUBLE
    (tests/jni_stub/fake_pgen.c, which decodes nothing) because a real handle needs a GPU."""
    out = str(tmp_path / "libshim_with_double.so")
    cmd = ["gcc", "-shared", "-fPIC", "-std=gnu99", "-Wall", "-Werror", "-Wno-unused-parameter", "-I", STUB, "-I", os.path.join(ROOT, "include"),
           SHIM, os.path.join(STUB, "mock_jvm.c"), os.path.join(STUB, "fake_pgen.c"), "-o", out]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    L = C.CDLL(out)
    for f in ("mock_env", "mock_new_string", "mock_new_bytes", "mock_element", "mock_data"):
        getattr(L, f).restype = C.c_void_p
    L.mock_new_string.argtypes = [C.c_char_p]
    L.mock_new_bytes.argtypes = [C.c_char_p, C.c_int]
    L.mock_element.argtypes = [C.c_void_p, C.c_int]
    L.mock_data.argtypes = L.mock_length.argtypes = L.mock_borrowed.argtypes = [C.c_void_p]
    L.mock_exception_message.restype = C.c_char_p
    create, submit, poll = (getattr(L, PKG + n) for n in ("create", "submit", "poll"))
    create.argtypes, create.restype = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32], C.c_int64
    submit.argtypes, submit.restype = [C.c_void_p, C.c_void_p, C.c_int64] + [C.c_void_p] * 3, None
    poll.argtypes, poll.restype = [C.c_void_p, C.c_void_p, C.c_int64, C.c_uint8], C.c_void_p
    env = L.mock_env()
    h = create(env, None, L.mock_new_string(b"p --mode=decode"), 0)
    assert h != 0
    rows = [(b"u1", b"article one .", b"ref 1"), (b"u2", b"", b"r\x00with nul"), (b"u3", b"caf\xc3\xa9", b"")]
    for u, a, r in rows:
        args = [L.mock_new_bytes(x, len(x)) for x in (u, a, r)]
        submit(env, None, h, *args)
        assert all(L.mock_borrowed(x) == 0 for x in args)                  # every GetByteArrayElements was released
    assert L.mock_length(poll(env, None, h, 0)) == 0                       # nothing finished yet: an empty array
    got = poll(env, None, h, 1)
    assert L.mock_length(got) == len(rows)
    for i, (u, a, r) in enumerate(rows):
        cols = L.mock_element(got, i)
        assert L.mock_length(cols) == 4
        vals = []
        for j in range(4):
            b = L.mock_element(cols, j)
            vals.append(C.string_at(L.mock_data(b), L.mock_length(b)))
        assert vals == [u, a, b"summary of " + u, r]
    assert L.mock_local_refs_deleted() == len(rows)
    for _ in range(13):
        submit(env, None, h, *[L.mock_new_bytes(b"x", 1) for _ in range(3)])
    assert L.mock_exception_message() is None
    submit(env, None, h, *[L.mock_new_bytes(b"x", 1) for _ in range(3)])   # the double holds 16 rows: the 17th fails
    assert L.mock_exception_message() == b"fake: full"
    getattr(L, PKG + "destroy").argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
    getattr(L, PKG + "destroy")(env, None, h)
