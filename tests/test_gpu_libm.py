"""csrc/hj_libm.h on the device against the same header compiled for the host (oracle/libhjo.so): every result must be the
SAME BITS — this is what lets the per-path parity tests compare kernels and restatement exactly."""
import numpy as np
import pytest

import henjou_renderer_b200 as hj
import oracle_binding as ob

pytestmark = pytest.mark.gpu


def _bits_equal(a, b):
    return (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))


def test_device_and_host_libm_give_the_same_bits():
    r = hj.Renderer(0)
    rng = np.random.default_rng(21)
    n = 1 << 20
    inf, nan = np.inf, np.nan
    special = np.array([0.0, -0.0, 1.0, -1.0, 0.5, -0.5, 2.0, -2.0, 3.0, 5.0, inf, -inf, nan, 1e-40, -1e-40, 1e-38, 3e38, -3e38, 0.9999, -0.9999,
                        1.0000001, 16777216.0, 16777217.0, 88.7, -103.9, 1e9, -1e9, 1e20], np.float32)
    # every float pattern class: random bit patterns cover denormals, huge values, NaN payloads
    anybits = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32).view(np.float32)
    unary = {
        "sinf": [rng.uniform(-7, 7, n), rng.uniform(-1e5, 1e5, n), anybits, special],
        "cosf": [rng.uniform(-7, 7, n), rng.uniform(-1e5, 1e5, n), anybits, special],
        "acosf": [rng.uniform(-1, 1, n), 1 - rng.uniform(0, 1e-3, n), anybits, special],
        "logf": [rng.uniform(0, 4, n), np.exp(rng.uniform(-87, 88, n)), anybits, special],
        "expf": [rng.uniform(-110, 90, n), anybits, special],
    }
    for fn, sets in unary.items():
        for a in sets:
            a = np.asarray(a, np.float32)
            got, want = r.libm(fn, a), ob.libm("restatement", fn, a)
            ok = _bits_equal(got, want)
            assert ok.all(), (fn, a[~ok][:5], got[~ok][:5], want[~ok][:5])
    X, Y = [g.ravel().astype(np.float32) for g in np.meshgrid(special, special)]
    binary = {
        "powf": [(rng.uniform(0, 1, n), rng.uniform(0, 8, n)), (rng.uniform(-3, 50, n), rng.uniform(-40, 40, n)), (rng.uniform(0, 1.2, n), np.full(n, 5.0)),
                 (rng.uniform(0, 4, n), np.full(n, 1 / 2.4)), (anybits, anybits[::-1].copy()), (X, Y)],
        "atan2f": [(rng.normal(size=n), rng.normal(size=n)), (anybits, anybits[::-1].copy()), (X, Y)],
    }
    for fn, sets in binary.items():
        for a, b in sets:
            a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
            got, want = r.libm(fn, a, b), ob.libm("restatement", fn, a, b)
            ok = _bits_equal(got, want)
            assert ok.all(), (fn, a[~ok][:5], b[~ok][:5], got[~ok][:5], want[~ok][:5])
    r.close()
