"""Restatement (oracle/libhjo.so) vs the reference's own headers compiled on the host (oracle/_ref/libhjref.so):
whole frames of the reference's cornell scenes must agree. Skipped where /root/reference is absent and
oracle/_ref was not prebuilt. Also: CPU BVH == brute force (the closest-hit definition)."""
import numpy as np
import pytest

import henjou_renderer_b200 as hj
from henjou_renderer_b200 import abi
import oracle_binding as ob
from oracle_binding import Oracle


def _pair(scene, fparams, xf, inv):
    if ob.lib("reference") is None:
        pytest.skip("oracle/_ref not built")
    a = Oracle(scene, "restatement", xforms=xf, inv_xforms=inv).render(fparams, debug=True)
    b = Oracle(scene, "reference", xforms=xf, inv_xforms=inv).render(fparams, debug=True)
    return a, b


@pytest.mark.parametrize("integrator", [abi.INTEGRATOR_PT, abi.INTEGRATOR_NEE, abi.INTEGRATOR_MIS])
def test_cornell_frame_matches_reference_headers(cornell, integrator):
    cornell.set_sky((0.8, 0.8, 0.8))
    xf, inv = cornell.transforms(1 / 24)
    fparams = hj.frame_params(cornell, 48, 48, 3, frame=1, max_depth=10, integrator=integrator)
    a, b = _pair(cornell, fparams, xf, inv)
    # same expression trees, same compiler flags: per-path radiance agrees to float rounding
    np.testing.assert_allclose(a["path_radiance"][..., :3], b["path_radiance"][..., :3], rtol=1e-5, atol=1e-7)
    assert np.array_equal(a["path_ids"][..., 0], b["path_ids"][..., 0])       # first-hit primitive ids
    assert a["rays_extend"] == b["rays_extend"] and a["rays_shadow"] == b["rays_shadow"]
    assert np.array_equal(a["rgba8"], b["rgba8"])


def test_textured_cornell_matches_reference_headers(cornell_tex):
    cornell_tex.set_sky((0.8, 0.8, 0.8))
    xf, inv = cornell_tex.transforms(0.0)
    cam = dict(position=np.array([4.5, 0.2, 0.1], np.float32), direction=np.array([-1, 0, 0], np.float32), fov=0.48)
    fparams = hj.frame_params(cornell_tex, 40, 40, 2, frame=0, max_depth=10, camera=cam)
    a, b = _pair(cornell_tex, fparams, xf, inv)
    np.testing.assert_allclose(a["path_radiance"][..., :3], b["path_radiance"][..., :3], rtol=1e-5, atol=1e-7)


def test_cpu_bvh_equals_brute_force(cornell):
    xf, inv = cornell.transforms(1 / 24)
    rng = np.random.default_rng(7)
    n = 4000
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3] = rng.uniform(-0.9, 0.9, (n, 3)) + np.array([0, 1.0, 0])
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays[:, 4:7] = d
    rays[:, 3] = 1e-3
    rays[:, 7] = 1e16
    o1 = Oracle(cornell, use_bvh=True, xforms=xf, inv_xforms=inv)
    o2 = Oracle(cornell, use_bvh=False, xforms=xf, inv_xforms=inv)
    p1, t1 = o1.trace(rays)
    p2, t2 = o2.trace(rays)
    assert np.array_equal(p1, p2) and np.array_equal(t1, t2)
    a1, _ = o1.trace(rays, any_hit=True)
    a2, _ = o2.trace(rays, any_hit=True)
    assert np.array_equal(a1, a2)
    assert (p1 != 0xffffffff).mean() > 0.5
