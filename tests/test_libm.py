"""henjou-renderer_b200/csrc/hj_libm.h — the elementary functions shared by the CUDA kernels and the CPU oracle — on the
host: accuracy against mpmath (true values) and glibc, C99 special cases. The device side of the same header is compared
bit for bit in tests/test_gpu_libm.py."""
import numpy as np
import pytest

import oracle_binding as ob

mp = pytest.importorskip("mpmath")
mp.mp.dps = 40


def ulp_err(got, true_mp):
    """|got - true| in ulps of float32 at the true value."""
    out = []
    for g, t in zip(got, true_mp):
        if t == 0:
            out.append(0.0 if g == 0 else np.inf)
            continue
        e = int(mp.floor(mp.log(abs(t), 2)))
        ulp = mp.mpf(2) ** (max(e, -126) - 23)
        out.append(float(abs(mp.mpf(float(g)) - t) / ulp))
    return np.array(out)


RNG = np.random.default_rng(7)
CASES = {
    "sinf": (np.concatenate([RNG.uniform(-2 * np.pi, 2 * np.pi, 3000), RNG.uniform(-1e4, 1e4, 1000), RNG.uniform(-1e-3, 1e-3, 200)]), None, mp.sin),
    "cosf": (np.concatenate([RNG.uniform(-2 * np.pi, 2 * np.pi, 3000), RNG.uniform(-1e4, 1e4, 1000), RNG.uniform(-1e-3, 1e-3, 200)]), None, mp.cos),
    "acosf": (np.concatenate([RNG.uniform(-1, 1, 3000), 1 - RNG.uniform(0, 1e-4, 300), -1 + RNG.uniform(0, 1e-4, 300)]), None, mp.acos),
    "logf": (np.concatenate([RNG.uniform(1e-6, 4, 3000), np.exp(RNG.uniform(-80, 80, 1000)), 1 + RNG.uniform(-1e-3, 1e-3, 300)]), None, mp.log),
    "expf": (np.concatenate([RNG.uniform(-20, 20, 3000), RNG.uniform(-87, 88, 1000)]), None, mp.exp),
    "powf": (np.concatenate([RNG.uniform(0, 1, 3000), RNG.uniform(0, 50, 1000)]), np.concatenate([RNG.uniform(0, 8, 3000), RNG.uniform(-3, 3, 1000)]), None),
    "atan2f": (RNG.normal(size=4000), RNG.normal(size=4000), None),
}


@pytest.mark.parametrize("fn", sorted(CASES))
def test_accuracy_against_mpmath_and_glibc(fn):
    a, b, f1 = CASES[fn]
    a = a.astype(np.float32)
    b = None if b is None else b.astype(np.float32)
    got = ob.libm("restatement", fn, a, b)
    glibc = ob.libm("glibc", fn, a, b)
    if f1 is not None:
        true = [f1(mp.mpf(float(x))) for x in a]
    elif fn == "powf":
        true = [mp.power(mp.mpf(float(x)), mp.mpf(float(y))) for x, y in zip(a, b)]
    else:
        true = [mp.atan2(mp.mpf(float(y)), mp.mpf(float(x))) for y, x in zip(a, b)]
    ok = np.isfinite(got) & np.isfinite(glibc) & (np.abs(glibc) > 1e-37)
    err = ulp_err(got[ok], [t for t, k in zip(true, ok) if k])
    # a double-precision evaluation rounded once to float: at most half an ulp plus the double's own error
    assert err.max() <= 0.5 + 1e-6, (fn, err.max())
    # glibc is within 1 ulp of the true value too, so the two agree to 1 ulp
    ulps = np.abs(got[ok].view(np.int32).astype(np.int64) - glibc[ok].view(np.int32).astype(np.int64))
    assert ulps.max() <= 1, (fn, ulps.max())
    assert (ulps == 0).mean() > 0.8       # glibc itself is within 1 ulp, not correctly rounded (atan2f: 84 % identical)


def test_special_cases_follow_c99():
    inf, nan = np.float32(np.inf), np.float32(np.nan)
    xs = np.array([0.0, -0.0, 1.0, -1.0, 2.0, -2.0, 0.5, -0.5, inf, -inf, nan, 1e-40, -8.0, 3.0], np.float32)
    ys = np.array([0.0, -0.0, 1.0, -1.0, 2.0, 3.0, -3.0, 0.5, -0.5, 5.0, inf, -inf, nan, 16777216.0, 16777217.0, 2.5], np.float32)
    X, Y = [g.ravel() for g in np.meshgrid(xs, ys)]
    got, want = ob.libm("restatement", "powf", X, Y), ob.libm("glibc", "powf", X, Y)
    same = (got.view(np.uint32) == want.view(np.uint32)) | (np.isnan(got) & np.isnan(want))
    close = np.isclose(got, want, rtol=2e-7, atol=0) & (np.signbit(got) == np.signbit(want))
    assert (same | close).all(), list(zip(X[~(same | close)], Y[~(same | close)], got[~(same | close)], want[~(same | close)]))
    for fn, arr in (("sinf", [0.0, -0.0, inf, -inf, nan]), ("cosf", [0.0, inf, nan]), ("acosf", [1.0, -1.0, 1.0000001, -1.5, nan, 0.0]),
                    ("logf", [0.0, -0.0, -1.0, inf, nan, 1.0, 1e-45]), ("expf", [-inf, inf, nan, 0.0, 89.0, -104.0, -90.0])):
        a = np.array(arr, np.float32)
        got, want = ob.libm("restatement", fn, a), ob.libm("glibc", fn, a)
        ok = (got.view(np.uint32) == want.view(np.uint32)) | (np.isnan(got) & np.isnan(want)) | (np.isclose(got, want, rtol=2e-7, atol=1e-45) & np.isfinite(want))
        assert ok.all(), (fn, a[~ok], got[~ok], want[~ok])
    ya = np.array([0.0, -0.0, 0.0, -0.0, 1.0, -1.0, inf, inf, -inf, 1.0, nan], np.float32)
    xa = np.array([1.0, 1.0, -1.0, -1.0, 0.0, 0.0, inf, -inf, -inf, -inf, 1.0], np.float32)
    got, want = ob.libm("restatement", "atan2f", ya, xa), ob.libm("glibc", "atan2f", ya, xa)
    ok = (got.view(np.uint32) == want.view(np.uint32)) | (np.isnan(got) & np.isnan(want)) | np.isclose(got, want, rtol=2e-7, atol=0)
    assert ok.all(), (ya[~ok], xa[~ok], got[~ok], want[~ok])
