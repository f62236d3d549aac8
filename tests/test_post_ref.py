"""CPU checks of the post-stage restatement (oracle/post_ref.py) the GPU tests compare hj_post against: the filter's
defining properties and float4ConvertColor's known answers (renderer.h:73-101)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import post_ref                                        # noqa: E402


def planes(H, W, seed=0):
    rng = np.random.default_rng(seed)
    return [rng.random((H, W, 4), dtype=np.float32) for _ in range(3)]


def test_constant_image_is_a_fixed_point():
    _, a, n = planes(23, 37)
    c = np.full((23, 37, 4), 0.25, np.float32)
    out = post_ref.denoise(c, a, n)
    np.testing.assert_allclose(out, c, rtol=2e-7)


def test_guides_stop_the_filter_at_edges():
    H, W = 32, 32
    c = np.zeros((H, W, 4), np.float32); c[:, W // 2:, :3] = 1.0; c[..., 3] = 1.0
    rng = np.random.default_rng(1)
    c[..., :3] += rng.normal(0, 0.05, (H, W, 3)).astype(np.float32)
    alb = np.zeros((H, W, 4), np.float32); alb[:, W // 2:, :3] = 1.0
    nrm = np.zeros((H, W, 4), np.float32); nrm[..., 1] = 1.0
    out = post_ref.denoise(c, alb, nrm)
    assert abs(out[:, :W // 2, :3].mean()) < 0.02 and abs(out[:, W // 2:, :3].mean() - 1.0) < 0.02      # no bleeding over the edge
    assert out[:, :W // 2, :3].std() < 0.5 * c[:, :W // 2, :3].std()                                   # noise reduced on each side
    assert np.array_equal(out[..., 3], c[..., 3])


def test_non_finite_samples_do_not_spread():
    c, a, n = planes(40, 40, 3)
    a[:] = 0.5; n[:] = 0.0
    c[20, 20, 0] = np.nan; c[5, 7, 1] = np.inf
    out = post_ref.denoise(c, a, n)
    assert np.isfinite(out).all()
    clean = c.copy(); clean[20, 20, 0] = 0.5; clean[5, 7, 1] = 0.5
    far = np.ones((40, 40), bool); far[20, 20] = far[5, 7] = False
    assert np.abs(out - post_ref.denoise(clean, a, n))[far].max() < 0.1


def test_upscale2x_reproduces_a_ramp_and_clamps():
    H, W = 5, 7
    img = np.zeros((H, W, 4), np.float32)
    img[..., 0] = np.arange(W, dtype=np.float32)[None, :]
    img[..., 1] = np.arange(H, dtype=np.float32)[:, None]
    up = post_ref.upscale2x(img)
    assert up.shape == (2 * H, 2 * W, 4)
    want_x = np.clip((np.arange(2 * W, dtype=np.float32) + 0.5) * 0.5 - 0.5, 0, W - 1)
    want_y = np.clip((np.arange(2 * H, dtype=np.float32) + 0.5) * 0.5 - 0.5, 0, H - 1)
    np.testing.assert_allclose(up[3, :, 0], want_x, atol=1e-6)
    np.testing.assert_allclose(up[:, 4, 1], want_y, atol=1e-6)


def test_rgba8_known_answers():
    c = np.array([[[0.0, 1.0, 0.5, 1.0], [0.002, 2.0, -1.0, 1.0]]], np.float32)
    q = post_ref.to_rgba8(c)
    assert q[0, 0].tolist() == [0, 255, 188, 255]          # 1.055 * 0.5^(1/2.4) - 0.055 = 0.7354 -> uint(188.26)
    assert q[0, 1].tolist() == [6, 255, 0, 255]            # linear branch 12.92 * 0.002 * 256 = 6.6; clamped above and below
