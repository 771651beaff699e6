"""pgen_attention_static -- the allocation-free twin of pgen_attention.

Some hosts of this library may not allocate device memory at all: the process' GPU memory is owned by
another allocator or capped by a fixed budget, or the calling loop cannot take the device-wide stall of
cudaMalloc/cudaFree.  For them the library keeps the attention scratch in a fixed 1 MiB device array.
This test drives both entry points through the C ABI on the same device buffers and requires the same
answers -- and the oracle's (attention_decoder.py:79-129) -- twice in a row with different shapes so that
whatever the first call leaves in the shared array is seen (and must be harmless) by the second; it also
checks that an over-sized request is refused.

Collected last on purpose (file name): a run that stops at the first failure has by then run every
other test.  It needs nothing from the other tests and also passes alone in a fresh process."""
import numpy as np
import pytest

from conftest import small_flags
from test_gpu_kernels import oracle_attention_from_features
from textsummarization_on_flink_b200 import native

pytestmark = pytest.mark.gpu

ATOL, RTOL = 5e-6, 1e-4          # the kernel-level tolerance of tests/test_gpu_kernels.py


def _case(torch, n, K, T, A, seed):
    g = np.random.default_rng(seed)
    host = dict(lens=g.integers(1, T + 1, size=n).astype(np.int32),
                enc=(g.standard_normal((n, T, A)) * 0.3).astype(np.float32),
                feat=(g.standard_normal((n, T, A)) * 0.8).astype(np.float32),
                dfeat=g.standard_normal((n * K, A)).astype(np.float32),
                cov=g.random((n * K, T)).astype(np.float32))
    return host, {k: torch.from_numpy(v).cuda() for k, v in host.items()}


def _run(torch, dec, fn, n, K, T, A, dev):
    attn = torch.full((n * K, T), 7.0, device="cuda")        # poison: every valid entry must be overwritten
    ctx = torch.full((n * K, A), 7.0, device="cuda")
    dec.attention(n, dev["feat"].data_ptr(), dev["enc"].data_ptr(), dev["lens"].data_ptr(), dev["dfeat"].data_ptr(),
                  dev["cov"].data_ptr(), attn.data_ptr(), ctx.data_ptr(), fn=fn)
    return attn.cpu().numpy(), ctx.cpu().numpy()


def test_static_scratch_attention_matches_allocating_entry_point(small_cfg, small_model_dir):
    import torch
    K, T, A = 4, 24, 2 * small_cfg["H"]
    with native.Decoder(small_flags(small_model_dir, small_cfg, beam=K, coverage=1, max_enc=T)) as dec:
        for call, (n, seed) in enumerate([(9, 1), (5, 2)]):          # two calls in a row, second one smaller
            host, dev = _case(torch, n, K, T, A, seed)
            ref_attn, ref_ctx = _run(torch, dec, "pgen_attention", n, K, T, A, dev)
            got_attn, got_ctx = _run(torch, dec, "pgen_attention_static", n, K, T, A, dev)
            for b in range(n):
                L = int(host["lens"][b])
                rows = slice(b * K, (b + 1) * K)
                np.testing.assert_allclose(got_attn[rows, :L], ref_attn[rows, :L], atol=ATOL, rtol=RTOL,
                                           err_msg="call %d article %d" % (call, b))
                np.testing.assert_allclose(got_attn[rows, :L].sum(axis=1), 1.0, atol=1e-5)
            np.testing.assert_allclose(got_ctx, ref_ctx, atol=ATOL, rtol=RTOL, err_msg="call %d" % call)
            # and against the oracle, on every entry: positions past an article's end hold exactly 0
            # (the poison written into the output buffers must be gone)
            ora_attn, ora_ctx = oracle_attention_from_features(small_cfg["W"], host, K, True)
            np.testing.assert_allclose(got_attn, ora_attn, atol=ATOL, rtol=RTOL, err_msg="call %d vs oracle" % call)
            np.testing.assert_allclose(got_ctx, ora_ctx, atol=ATOL, rtol=RTOL, err_msg="call %d vs oracle" % call)
            for b in range(n):
                assert not got_attn[b * K:(b + 1) * K, int(host["lens"][b]):].any()


def test_static_scratch_rejects_requests_beyond_its_array(small_cfg, small_model_dir):
    import torch
    K, T, A = 4, 24, 2 * small_cfg["H"]
    per_article = 4 * K * T
    n_big = (1 << 20) // per_article + 8                               # just past the 1 MiB array
    with native.Decoder(small_flags(small_model_dir, small_cfg, beam=K, coverage=1, max_enc=T)) as dec:
        host, dev = _case(torch, n_big, K, T, A, 3)
        attn = torch.zeros(n_big * K, T, device="cuda")
        ctx = torch.zeros(n_big * K, A, device="cuda")
        args = (n_big, dev["feat"].data_ptr(), dev["enc"].data_ptr(), dev["lens"].data_ptr(), dev["dfeat"].data_ptr(),
                dev["cov"].data_ptr(), attn.data_ptr(), ctx.data_ptr())
        with pytest.raises(native.PgenError) as e:
            dec.attention(*args, fn="pgen_attention_static")
        assert e.value.code == 1 and "static array holds 1048576" in str(e.value)
        assert not attn.any().item()                                   # refused before any GPU work
        dec.attention(*args, fn="pgen_attention")                      # the allocating entry point still takes it
        assert attn.any().item()
