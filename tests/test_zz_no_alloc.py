"""hj_trace_noalloc — the batch ray query for callers that cannot call cudaMalloc (no
counterpart in the reference). Its device scratch is one fixed `static __device__` array inside the library instead of buffers the
context grows. Through the C ABI it must give the SAME BITS as hj_trace — (prim, t, b1, b2) are exact in every trace test of this
suite — on two calls in a row (the array keeps the previous call's contents: the second call is shorter and uses other rays, so a
stale cursor or stale results would show), for closest-hit and any-hit queries, and refuse batches beyond its stated maximum.
This file sorts last on purpose (`pytest --collect-only`)."""
import os

import numpy as np
import pytest

import henjou_renderer_b200 as hj
from henjou_renderer_b200 import abi

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _rays(n, seed):
    rng = np.random.default_rng(seed)
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3] = rng.uniform((-0.95, 0.0, -0.95), (0.95, 2.2, 0.95), (n, 3))
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays[:, 4:7] = d
    rays[:, 3] = 1e-3
    rays[:, 7] = 1e16
    return rays


def test_trace_noalloc_equals_trace_on_two_calls_in_a_row():
    scene = hj.load_npz(os.path.join(ROOT, "tests", "golden", "cornelbox.npz"))
    xf, inv = scene.transforms(1 / 24)
    r = hj.Renderer(0)
    try:
        r.upload(scene); r.build_accel(xf, inv)
        first, second = _rays(abi.TRACE_NOALLOC_MAX_RAYS, 31), _rays(20001, 32)      # the maximum, then a shorter, ragged batch
        want = [r.trace(first), r.trace(second), r.trace(second, any_hit=True)]      # existing entry point (grows context buffers)
        got1 = r.trace_noalloc(first)
        got2 = r.trace_noalloc(second)                                               # straight after: the array still holds call 1
        got3 = r.trace_noalloc(second, any_hit=True)
        for (p, t), (p0, t0) in zip((got1, got2), want[:2]):
            assert np.array_equal(p, p0)
            assert np.array_equal(t.view(np.uint32), t0.view(np.uint32))            # t, b1, b2: the same bits
        assert np.array_equal(got3[0], want[2][0])
        assert (want[0][0] != 0xffffffff).mean() > 0.5 and (want[2][0] == 1).any()   # the rays really hit things
        # beyond the stated maximum: a clear error, nothing traced
        with pytest.raises(hj.HenjouError, match="HJ_TRACE_NOALLOC_MAX_RAYS"):
            r.trace_noalloc(_rays(abi.TRACE_NOALLOC_MAX_RAYS + 1, 33))
        # a pending asynchronous build is refused (completing it may allocate); after hj_accel_wait the call works again
        r.build_accel(xf, inv, sync=False)
        with pytest.raises(hj.HenjouError, match="hj_accel_wait"):
            r.trace_noalloc(second)
        r.accel_wait()
        p, t = r.trace_noalloc(second)
        assert np.array_equal(p, want[1][0]) and np.array_equal(t.view(np.uint32), want[1][1].view(np.uint32))
        # an empty batch is fine
        p, t = r.trace_noalloc(np.zeros((0, 8), np.float32))
        assert p.size == 0
    finally:
        r.close()
    # without an acceleration structure: an error, not a fault
    r2 = hj.Renderer(0)
    try:
        r2.upload(scene)
        with pytest.raises(hj.HenjouError, match="hj_accel_build"):
            r2.trace_noalloc(_rays(16, 1))
    finally:
        r2.close()
