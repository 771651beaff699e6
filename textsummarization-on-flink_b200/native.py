"""ctypes binding of libpgen_cuda.so (include/pgen.h).  No fallback: importing works anywhere, but
every call that needs the GPU raises PgenError when the library or the device is missing."""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PGEN_LIB") or os.path.join(_HERE, "libpgen_cuda.so")   # PGEN_LIB: a diagnostic build (csrc/build.sh)

EXPORTS = ["pgen_create", "pgen_submit", "pgen_poll", "pgen_rows_free", "pgen_stats", "pgen_last_error",
           "pgen_destroy", "pgen_flag_int", "pgen_text_encode", "pgen_text_summary", "pgen_decode_ids",
           "pgen_stage_upload", "pgen_stage_run", "pgen_stage_sync", "pgen_stage_download", "pgen_stream",
           "pgen_debug_fetch", "pgen_set_option", "pgen_attention",
           "pgen_attention_static", "pgen_debug_vocab_logits", "pgen_stage_slot",
           "pgen_example_parse", "pgen_example_write", "pgen_submit_example", "pgen_device_count",
           "pgen_debug_where", "pgen_text_open", "pgen_text_vocab_size", "pgen_text_encode_article",
           "pgen_text_decode_summary", "pgen_text_close", "pgen_parse_flag"]

ERR_NAMES = {1: "PGEN_ERR_ARG", 2: "PGEN_ERR_IO", 3: "PGEN_ERR_CUDA", 4: "PGEN_ERR_STATE", 5: "PGEN_ERR_UNSUPPORTED"}


class PgenError(RuntimeError):
    """Mirrors the reference wrapping every failure in RuntimeException (TFModel.java:78-80)."""

    def __init__(self, code: int, msg: str):
        super().__init__("%s: %s" % (ERR_NAMES.get(code, code), msg))
        self.code = code


class Row(C.Structure):
    _fields_ = [("uuid", C.c_void_p), ("uuid_len", C.c_size_t), ("article", C.c_void_p), ("article_len", C.c_size_t),
                ("summary", C.c_void_p), ("summary_len", C.c_size_t), ("reference", C.c_void_p), ("reference_len", C.c_size_t)]


class Rows(C.Structure):
    _fields_ = [("rows", C.POINTER(Row)), ("n_rows", C.c_size_t), ("owner_", C.c_void_p)]


class Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("rows_in", "rows_out", "batches", "tokens_in", "tokens_out", "decode_steps",
                                           "article_steps", "early_stops", "kernel_launches")] + \
               [(n, C.c_double) for n in ("ms_host_text", "ms_h2d", "ms_encoder", "ms_decoder", "ms_d2h",
                                          "ms_attention", "ms_vocab", "ms_select", "prof_steps")] + \
               [(n, C.c_uint64) for n in ("selection_widened", "selection_fallbacks")] + \
               [(n, C.c_double) for n in ("ms_enc_recurrence", "ms_enc_features")]


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PgenError(2, "libpgen_cuda.so is not built (%s); run __graft_entry__.build()" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        i32p, f32p, f64p = C.POINTER(C.c_int32), C.POINTER(C.c_float), C.POINTER(C.c_double)
        L.pgen_create.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]
        L.pgen_submit.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
        L.pgen_poll.argtypes = [C.c_void_p, C.c_int, C.POINTER(Rows)]
        L.pgen_rows_free.argtypes = [C.POINTER(Rows)]
        L.pgen_rows_free.restype = None
        L.pgen_stats.argtypes = [C.c_void_p, C.POINTER(Stats)]
        L.pgen_last_error.argtypes = [C.c_void_p]
        L.pgen_last_error.restype = C.c_char_p
        L.pgen_destroy.argtypes = [C.c_void_p]
        L.pgen_destroy.restype = None
        L.pgen_flag_int.argtypes = [C.c_void_p, C.c_char_p]
        L.pgen_flag_int.restype = C.c_int64
        L.pgen_text_encode.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, i32p, i32p, C.c_int, i32p, i32p]
        L.pgen_text_summary.argtypes = [C.c_void_p, i32p, C.c_int, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.pgen_decode_ids.argtypes = [C.c_void_p, C.c_int, i32p, i32p, C.c_int, i32p, i32p, i32p, i32p, f32p, f64p]
        L.pgen_stage_upload.argtypes = [C.c_void_p, C.c_int, i32p, i32p, C.c_int, i32p, i32p]
        L.pgen_stage_run.argtypes = [C.c_void_p]
        L.pgen_stage_sync.argtypes = [C.c_void_p]
        L.pgen_stage_download.argtypes = [C.c_void_p, i32p, i32p, f32p, f64p]
        L.pgen_stream.argtypes = [C.c_void_p]
        L.pgen_stage_slot.argtypes = [C.c_void_p, C.c_int]
        L.pgen_example_parse.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(Row)]
        L.pgen_example_write.argtypes = [C.POINTER(Row), C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.pgen_submit_example.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
        L.pgen_stream.restype = C.c_void_p
        L.pgen_device_count.argtypes = []
        L.pgen_device_count.restype = C.c_int
        L.pgen_debug_fetch.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.pgen_debug_where.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.pgen_set_option.argtypes = [C.c_void_p, C.c_char_p]
        L.pgen_attention.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 7
        L.pgen_debug_vocab_logits.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.pgen_attention_static.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 7
        L.pgen_text_open.argtypes = [C.c_char_p, C.c_int64, C.c_int, C.POINTER(C.c_void_p)]
        L.pgen_text_vocab_size.argtypes = [C.c_void_p]
        L.pgen_text_encode_article.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, i32p, i32p, C.c_int, i32p, i32p]
        L.pgen_text_decode_summary.argtypes = [C.c_void_p, i32p, C.c_int, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.pgen_text_close.argtypes = [C.c_void_p]
        L.pgen_text_close.restype = None
        L.pgen_parse_flag.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(C.c_int64)]
        _lib = L
    return _lib


def parse_flag(hyperparams: str, name: str) -> int:
    """Value of an integer / boolean flag as pgen_create would parse it (no device needed)."""
    v = C.c_int64()
    rc = lib().pgen_parse_flag(hyperparams.encode(), name.encode(), C.byref(v))
    if rc != 0:
        raise PgenError(rc, (lib().pgen_last_error(None) or b"").decode(errors="replace"))
    return int(v.value)


def device_count() -> int:
    """CUDA devices visible to the library (0 without a GPU)."""
    return int(lib().pgen_device_count())


def _p(a: np.ndarray, t):
    return a.ctypes.data_as(C.POINTER(t))


def pack_articles(arts: Sequence[Tuple[np.ndarray, np.ndarray, int]], ld: int):
    """[(enc_ids, ext_ids, n_oov)] -> padded int32 arrays for pgen_decode_ids."""
    n = len(arts)
    ids = np.zeros((n, ld), np.int32)
    ext = np.zeros((n, ld), np.int32)
    lens = np.zeros((n,), np.int32)
    noov = np.zeros((n,), np.int32)
    for i, (a, e, k) in enumerate(arts):
        lens[i] = len(a)
        noov[i] = k
        ids[i, :len(a)] = a
        ext[i, :len(a)] = e
    return ids, ext, lens, noov


def example_parse(buf: bytes):
    """tf.Example bytes -> (uuid, article, summary, reference) via the library's codec (no GPU needed)."""
    r = Row()
    rc = lib().pgen_example_parse(buf, len(buf), C.byref(r))
    if rc != 0:
        raise PgenError(rc, (lib().pgen_last_error(None) or b"").decode(errors="replace"))
    return tuple(C.string_at(getattr(r, f), getattr(r, f + "_len")) if getattr(r, f) else b""
                 for f in ("uuid", "article", "summary", "reference"))


def example_write(uuid: bytes, article: bytes, summary: bytes, reference: bytes) -> bytes:
    keep = [C.create_string_buffer(x, len(x)) for x in (uuid, article, summary, reference)]
    r = Row(C.cast(keep[0], C.c_void_p), len(uuid), C.cast(keep[1], C.c_void_p), len(article),
            C.cast(keep[2], C.c_void_p), len(summary), C.cast(keep[3], C.c_void_p), len(reference))
    n = C.c_size_t()
    lib().pgen_example_write(C.byref(r), None, 0, C.byref(n))
    out = C.create_string_buffer(max(1, n.value))
    rc = lib().pgen_example_write(C.byref(r), out, n.value, C.byref(n))
    if rc != 0:
        raise PgenError(rc, (lib().pgen_last_error(None) or b"").decode(errors="replace"))
    return out.raw[:n.value]


class TextPath:
    """The library's host text path (csrc/text.cc) on a handle of its own: needs no GPU."""

    def __init__(self, vocab_path: str, vocab_size: int, max_enc_steps: int = 400):
        self._L = lib()
        self._t = C.c_void_p()
        rc = self._L.pgen_text_open(vocab_path.encode(), vocab_size, max_enc_steps, C.byref(self._t))
        if rc != 0:
            raise PgenError(rc, (self._L.pgen_last_error(None) or b"").decode(errors="replace"))
        self.max_enc_steps = max_enc_steps

    def _ck(self, rc: int):
        if rc != 0:
            raise PgenError(rc, (self._L.pgen_last_error(None) or b"").decode(errors="replace"))

    def close(self):
        if self._t:
            self._L.pgen_text_close(self._t)
            self._t = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def vocab_size(self) -> int:
        return int(self._L.pgen_text_vocab_size(self._t))

    def encode(self, article: bytes):
        """-> (enc ids with OOV -> [UNK], extended ids, number of distinct in-article OOVs)"""
        cap = self.max_enc_steps
        ids = np.zeros(cap, np.int32)
        ext = np.zeros(cap, np.int32)
        n, k = C.c_int32(), C.c_int32()
        self._ck(self._L.pgen_text_encode_article(self._t, article, len(article), _p(ids, C.c_int32), _p(ext, C.c_int32), cap,
                                                   C.byref(n), C.byref(k)))
        return ids[:n.value].copy(), ext[:n.value].copy(), k.value

    def summary(self, tokens: Sequence[int]) -> bytes:
        """summary string of decoded token ids, for the article last passed to encode()"""
        t = np.asarray(tokens, np.int32)
        ln = C.c_size_t()
        buf = C.create_string_buffer(64 + 64 * max(1, len(t)))
        self._ck(self._L.pgen_text_decode_summary(self._t, _p(t, C.c_int32), len(t), buf, len(buf), C.byref(ln)))
        if ln.value > len(buf):
            buf = C.create_string_buffer(ln.value)
            self._ck(self._L.pgen_text_decode_summary(self._t, _p(t, C.c_int32), len(t), buf, len(buf), C.byref(ln)))
        return buf.raw[:ln.value]


class Decoder:
    """One handle = one Flink subtask's model instance on one GPU."""

    def __init__(self, hyperparams: str, device: int = 0):
        self._L = lib()
        self._h = C.c_void_p()
        rc = self._L.pgen_create(hyperparams.encode(), device, C.byref(self._h))
        if rc != 0:
            raise PgenError(rc, (self._L.pgen_last_error(None) or b"").decode(errors="replace"))
        self.max_enc_steps = self.flag("max_enc_steps")
        self.max_dec_steps = self.flag("max_dec_steps")
        self.articles_per_batch = self.flag("articles_per_batch")
        self.beam_size = self.flag("beam_size")
        self._n_staged = 0

    def _ck(self, rc: int):
        if rc != 0:
            raise PgenError(rc, (self._L.pgen_last_error(self._h) or b"").decode(errors="replace"))

    def close(self):
        if self._h:
            self._L.pgen_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def flag(self, name: str) -> int:
        return int(self._L.pgen_flag_int(self._h, name.encode()))

    def set_option(self, kv: str):
        self._ck(self._L.pgen_set_option(self._h, kv.encode()))

    # ---- row level (the operator's open / flatMap / close pattern) ----
    def submit(self, uuid: bytes, article: bytes, reference: bytes):
        self._ck(self._L.pgen_submit(self._h, uuid, len(uuid), article, len(article), reference, len(reference)))

    def submit_example(self, example: bytes):
        self._ck(self._L.pgen_submit_example(self._h, example, len(example)))

    def poll(self, flush: bool = False) -> List[Tuple[bytes, bytes, bytes, bytes]]:
        rows = Rows()
        self._ck(self._L.pgen_poll(self._h, 1 if flush else 0, C.byref(rows)))
        out = []
        for i in range(rows.n_rows):
            r = rows.rows[i]
            out.append(tuple(C.string_at(getattr(r, f), getattr(r, f + "_len"))
                             for f in ("uuid", "article", "summary", "reference")))
        self._L.pgen_rows_free(C.byref(rows))
        return out

    def stats(self) -> dict:
        s = Stats()
        self._ck(self._L.pgen_stats(self._h, C.byref(s)))
        return {n: getattr(s, n) for n, _ in Stats._fields_}

    # ---- text path ----
    def text_encode(self, article: bytes):
        cap = self.max_enc_steps
        ids = np.zeros(cap, np.int32)
        ext = np.zeros(cap, np.int32)
        n, k = C.c_int32(), C.c_int32()
        self._ck(self._L.pgen_text_encode(self._h, article, len(article), _p(ids, C.c_int32), _p(ext, C.c_int32), cap,
                                           C.byref(n), C.byref(k)))
        return ids[:n.value].copy(), ext[:n.value].copy(), k.value

    def text_summary(self, tokens: Sequence[int]) -> bytes:
        t = np.asarray(tokens, np.int32)
        ln = C.c_size_t()
        buf = C.create_string_buffer(64 + 64 * max(1, len(t)))
        self._ck(self._L.pgen_text_summary(self._h, _p(t, C.c_int32), len(t), buf, len(buf), C.byref(ln)))
        if ln.value > len(buf):
            buf = C.create_string_buffer(ln.value)
            self._ck(self._L.pgen_text_summary(self._h, _p(t, C.c_int32), len(t), buf, len(buf), C.byref(ln)))
        return buf.raw[:ln.value]

    # ---- tensor level ----
    def _outs(self, n):
        S = self.max_dec_steps
        return (np.zeros((n, S), np.int32), np.zeros((n,), np.int32), np.zeros((n, S), np.float32),
                np.zeros((n,), np.float64))

    def decode_ids(self, ids, ext, lens, noov):
        n, ld = ids.shape
        tok, ln, lp, avg = self._outs(n)
        self._ck(self._L.pgen_decode_ids(self._h, n, _p(ids, C.c_int32), _p(ext, C.c_int32), ld, _p(lens, C.c_int32),
                                          _p(noov, C.c_int32), _p(tok, C.c_int32), _p(ln, C.c_int32),
                                          _p(lp, C.c_float), _p(avg, C.c_double)))
        return tok, ln, lp, avg

    def stage_upload(self, ids, ext, lens, noov):
        n, ld = ids.shape
        self._n_staged = n
        self._ck(self._L.pgen_stage_upload(self._h, n, _p(ids, C.c_int32), _p(ext, C.c_int32), ld,
                                            _p(lens, C.c_int32), _p(noov, C.c_int32)))

    def stage_slot(self, slot: int):
        self._ck(self._L.pgen_stage_slot(self._h, slot))

    def stage_run(self):
        self._ck(self._L.pgen_stage_run(self._h))

    def stage_sync(self):
        self._ck(self._L.pgen_stage_sync(self._h))

    def stage_download(self):
        tok, ln, lp, avg = self._outs(self._n_staged)
        self._ck(self._L.pgen_stage_download(self._h, _p(tok, C.c_int32), _p(ln, C.c_int32), _p(lp, C.c_float),
                                              _p(avg, C.c_double)))
        return tok, ln, lp, avg

    def attention(self, n_articles: int, enc_feat, enc_states, lens, dfeat, cov, attn, ctx, fn: str = "pgen_attention"):
        """Kernel-level attention pass on DEVICE pointers (ints, e.g. torch.Tensor.data_ptr())."""
        self._ck(getattr(self._L, fn)(self._h, n_articles, enc_feat, enc_states, lens, dfeat, cov, attn, ctx))

    def vocab_logits(self, rows: int, out_ptr: int, logits_ptr: int, use_tc: bool):
        self._ck(self._L.pgen_debug_vocab_logits(self._h, rows, out_ptr, logits_ptr, 1 if use_tc else 0))

    def stream_ptr(self) -> int:
        return int(self._L.pgen_stream(self._h) or 0)

    def debug_where(self) -> list:
        """Callable from another thread while a call on this handle is blocked (include/pgen.h: pgen_debug_where)."""
        out = np.zeros(128, np.int64)
        n = self._L.pgen_debug_where(self._h, out.ctypes.data_as(C.c_void_p), 128)
        return [int(x) for x in out[:max(0, n)]]

    def debug_fetch(self, name: str, count: int, dtype=np.float32) -> np.ndarray:
        out = np.zeros(count, dtype)
        n = C.c_size_t()
        self._ck(self._L.pgen_debug_fetch(self._h, name.encode(), out.ctypes.data_as(C.c_void_p), out.nbytes, C.byref(n)))
        return out[:n.value]
