"""GPU parity at BASELINE.json's FULL sizes, through properties that do not need a full CPU render: the C2 stand-in
(1.85 M triangles, 1920x1080) and the C5 stand-in (10.2 M triangles).
  * the structural BVH check (k_verify, on for the whole suite) passes on the full builds, for both builders;
  * closest-hit (prim, t, b1, b2) of random rays are bit-exact against the CPU oracle's BVH walk of the same scene;
  * the closest hit is a pure function of (ray, scene): frames are bit-identical whatever the tree (PLOC / LBVH, leaf
    split on / off, rebuild / refit) and however the frame is cut into passes (pixel bands, pipelines, pass size)."""
import numpy as np
import pytest

import henjou_renderer_b200 as hj
from henjou_renderer_b200 import abi, scenes
from oracle_binding import Oracle
from test_gpu_parity import random_rays

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    r = hj.Renderer(0)
    r.set_option("verify_accel", 1)
    yield r
    r.close()


@pytest.fixture(scope="module")
def c2():
    return scenes.rtcamp_like()


def _frame(gpu, fp):
    gpu.render(fp)
    return [gpu.readback(a).copy() for a in (abi.AOV_ACCUM, abi.AOV_ALBEDO, abi.AOV_NORMAL, abi.AOV_RGBA8)], gpu.stats()


def _same(a, b):
    return all(np.array_equal(x.view(np.uint8), y.view(np.uint8)) for x, y in zip(a, b))


def test_c2_full_rays_against_the_oracle_bvh(gpu, c2):
    assert c2.n_triangles > 1_800_000
    xf, inv = c2.transforms(1 / 24.0)
    gpu.upload(c2); gpu.build_accel(xf, inv)
    lo, hi = np.array([-14.0, 0.05, -14.0]), np.array([14.0, 9.0, 14.0])
    rays = random_rays(100_000, 11, lo=lo, hi=hi)
    want_p, want_t = Oracle(c2, xforms=xf, inv_xforms=inv).trace(rays)
    assert (want_p != 0xffffffff).mean() > 0.5
    for builder in (1, 0):
        gpu.set_option("builder", builder); gpu.build_accel(xf, inv)
        p, t = gpu.trace(rays)
        assert np.array_equal(p, want_p) and np.array_equal(t, want_t), builder
    gpu.set_option("builder", 1)
    gpu.build_accel(xf, inv, mode=abi.ACCEL_REFIT)
    p, t = gpu.trace(rays)
    assert np.array_equal(p, want_p) and np.array_equal(t, want_t)


def test_c2_full_frame_does_not_depend_on_tree_or_pass_layout(gpu, c2):
    xf, inv = c2.transforms(1 / 24.0)
    gpu.upload(c2); gpu.build_accel(xf, inv)
    fp = hj.frame_params(c2, 1920, 1080, 2, frame=1, max_depth=10, integrator=abi.INTEGRATOR_MIS)
    ref, st = _frame(gpu, fp)
    assert st["passes"] >= 2                              # 2 spp: one pass per pipeline (auto: two pipelines)
    assert np.isfinite(ref[0][..., :3]).mean() > 0.9999 and np.nansum(ref[0][..., :3]) > 0
    for key, val, back in (("bands", 1, 0), ("bands", 5, 0), ("pipelines", 1, 0), ("pipelines", 8, 0), ("paths_per_pass", 1 << 20, 0),
                           ("mis_emitter_first", 0, 1)):
        gpu.set_option(key, val)
        got, _ = _frame(gpu, fp)
        gpu.set_option(key, back)
        assert _same(ref, got), (key, val)
    for key, val, back in (("leaf_split", 0, 1), ("builder", 0, 1), ("leaf_max", 1, 3), ("ploc_tail", 0, 1)):
        gpu.set_option(key, val); gpu.build_accel(xf, inv)
        got, _ = _frame(gpu, fp)
        gpu.set_option(key, back)
        assert _same(ref, got), (key, val)
    gpu.build_accel(xf, inv)
    gpu.build_accel(xf, inv, mode=abi.ACCEL_REFIT)
    got, _ = _frame(gpu, fp)
    assert _same(ref, got)


def test_c5_full_build_verifies_and_builders_agree(gpu):
    s = scenes.icosphere_field()
    assert s.n_triangles > 10_000_000
    gpu.upload(s); gpu.build_accel()
    lo, hi = s.vertices.min(0), s.vertices.max(0)
    rays = random_rays(200_000, 12, lo=lo, hi=hi)
    p1, t1 = gpu.trace(rays)
    a1, _ = gpu.trace(rays, any_hit=True)
    assert (p1 != 0xffffffff).mean() > 0.5
    assert np.array_equal(a1 != 0xffffffff, p1 != 0xffffffff)          # an occlusion query finds a hit iff there is one
    gpu.set_option("builder", 0); gpu.build_accel()
    p0, t0 = gpu.trace(rays)
    gpu.set_option("builder", 1)
    assert np.array_equal(p0, p1) and np.array_equal(t0, t1)
    # a sample of the rays against the CPU oracle (brute force over 10 M triangles would take minutes: BVH walk)
    want_p, want_t = Oracle(s).trace(rays[:20_000])
    assert np.array_equal(p1[:20_000], want_p) and np.array_equal(t1[:20_000], want_t)
