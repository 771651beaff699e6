"""The error bound behind the exactness guard of the tensor-core candidate selection (csrc/vocab_tc.cu, step 5 of
vocab_final_kernel): a word's tf32 logit differs from its fp32 logit by at most

    eps = 1.1e-3 * |out|_2 * max_w |w_w|_2 + 4e-6 * (|M| + 1)

Checked here on the CPU by emulating what the tensor core is given: both operands rounded to tf32 with cvt.rna (round to
nearest, ties away, 10 explicit mantissa bits), products exact, sum in fp32 with truncation after every addition (the
worst rounding mode an accumulator can have).  Random and adversarial operands, hidden_dim 256."""
import numpy as np


def tf32_rna(x):
    b = np.asarray(x, np.float32).view(np.uint32).astype(np.uint64)
    b = (b + 0x1000) & 0xFFFFE000                      # add half an ulp of the 13 dropped bits, then drop them
    return b.astype(np.uint32).view(np.float32)


def trunc_sum(products):
    """fp32 accumulation with round-toward-zero after every addition"""
    acc = np.zeros(products.shape[0], np.float32)
    for k in range(products.shape[1]):
        s = acc.astype(np.float64) + products[:, k].astype(np.float64)
        f = s.astype(np.float32)
        over = np.abs(f.astype(np.float64)) > np.abs(s)
        f[over] = np.nextafter(f[over], np.float32(0))
        acc = f
    return acc


def test_tf32_rounding_emulation():
    x = np.float32(1.0) + np.float32(2.0 ** -11)       # exactly half way between two tf32 values: ties away from zero
    assert tf32_rna(x) == np.float32(1.0 + 2.0 ** -10) and tf32_rna(-x) == np.float32(-(1.0 + 2.0 ** -10))
    r = np.random.default_rng(0).standard_normal(10000).astype(np.float32)
    assert np.all(np.abs(tf32_rna(r) - r) <= np.abs(r) * 2.0 ** -11)
    assert np.all((tf32_rna(r).view(np.uint32) & 0x1FFF) == 0)


def test_guard_bound_holds():
    rng = np.random.default_rng(1)
    H, n = 256, 4000
    worst = 0.0
    for scale_o, scale_w, mode in [(1.0, 0.7, "normal"), (30.0, 0.7, "normal"), (1.0, 1e-4, "normal"), (3.0, 2.0, "sign"), (1.0, 1.0, "halfway")]:
        o = (rng.standard_normal(H) * scale_o).astype(np.float32)
        w = (rng.standard_normal((n, H)) * scale_w).astype(np.float32)
        if mode == "sign":                              # every product positive: no cancellation, the errors add up
            w = np.abs(w) * np.sign(o)
        if mode == "halfway":                           # mantissas just below a rounding boundary: the largest relative error
            o = (np.sign(o) * (1.0 + 2.0 ** -11 - 2.0 ** -23) * 2.0 ** rng.integers(-2, 3, H)).astype(np.float32)
            w = (np.sign(w) * (1.0 + 2.0 ** -11 - 2.0 ** -23) * 2.0 ** rng.integers(-2, 3, (n, H))).astype(np.float32)
        exact = (w.astype(np.float64) @ o.astype(np.float64))
        prods = tf32_rna(w).astype(np.float64) * tf32_rna(o).astype(np.float64)[None, :]
        assert np.array_equal(prods, prods.astype(np.float32).astype(np.float64))          # 11 x 11 significant bits: exact in fp32
        got = trunc_sum(prods.astype(np.float32)).astype(np.float64)
        M = float(np.max(exact))
        eps = 1.1e-3 * float(np.linalg.norm(o.astype(np.float64))) * float(np.max(np.linalg.norm(w.astype(np.float64), axis=1))) + 4e-6 * (abs(M) + 1.0)
        err = float(np.max(np.abs(got - exact)))
        assert err <= eps, (mode, err, eps)
        worst = max(worst, err / eps)
        print(mode, scale_o, scale_w, 'error / bound = %.3f' % (err / eps))
    assert 0.01 < worst <= 1.0                          # the bound is not vacuous either
