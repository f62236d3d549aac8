"""GPU parity on the procedural stand-ins of BASELINE.json's configs, at sizes the CPU oracle finishes in seconds:
thin-film LUT + EC-GGX metals + clearcoat/sheen (C2), negative-IOR glass at depth 32 (C3), animated instances with
BVH rebuild vs refit (C4), many-triangle traversal (C5), and the two BVH builders against each other."""
import numpy as np
import pytest

import henjou_renderer_b200 as hj
from henjou_renderer_b200 import abi, scenes
from oracle_binding import Oracle
from test_gpu_parity import path_mismatch, random_rays
import parity_tools as pt

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    r = hj.Renderer(0)
    yield r
    r.close()


def box_rays(scene, n, seed):
    lo, hi = scene.vertices.min(0), scene.vertices.max(0)
    return random_rays(n, seed, lo=lo - 0.1, hi=hi + 0.1)


@pytest.mark.parametrize("integrator", [abi.INTEGRATOR_MIS, abi.INTEGRATOR_NEE, abi.INTEGRATOR_PT])
def test_c2_thinfilm_metal_scene(gpu, integrator):
    s = scenes.rtcamp_like(statue_subdiv=4, detail=0.08)
    assert any(m["is_thinfilm"] for m in s.materials) and any(m["metallic"] > 0.5 for m in s.materials) and s.lut is not None
    fp = hj.frame_params(s, 112, 63, 3, frame=1, max_depth=10, integrator=integrator, debug_paths=1)
    gpu.upload(s); gpu.build_accel()
    gpu.render(fp)
    want = Oracle(s).render(fp, debug=True)
    ids = gpu.readback(abi.AOV_PATH_IDS); rad = gpu.readback(abi.AOV_PATH_RADIANCE)
    pt.compare_paths("C2 stand-in 112x63x3 depth 10 integrator %d" % integrator, rad, ids, want, exact=True)   # thin-film LUT, EC-GGX walk, clearcoat: same bits
    assert pt.bit_equal(gpu.readback(abi.AOV_COLOR), want["color"])
    assert pt.bit_equal(gpu.readback(abi.AOV_ALBEDO), want["albedo"])


def test_c3_negative_ior_depth32(gpu):
    s = scenes.metaglass_room(subdiv=3)
    assert any(m["ideal_specular"] for m in s.materials)
    fp = hj.frame_params(s, 80, 60, 2, frame=3, max_depth=32, integrator=abi.INTEGRATOR_MIS, debug_paths=1)
    gpu.upload(s); gpu.build_accel()
    gpu.render(fp)
    want = Oracle(s).render(fp, debug=True)
    ids = gpu.readback(abi.AOV_PATH_IDS); rad = gpu.readback(abi.AOV_PATH_RADIANCE)
    pt.compare_paths("C3 stand-in 80x60x2 depth 32", rad, ids, want, exact=True)     # up to 32 Fresnel / roulette decisions per path, all the same
    deep = want["path_radiance"][..., 3] > 20
    assert deep.any()                           # the scene really produces deep paths
    st = gpu.stats()
    assert int(st["rays_extend"]) == int(want["traced_extend"]) and int(st["rays_shadow"]) == int(want["traced_shadow"])


def test_c4_animation_rebuild_equals_refit(gpu):
    s = scenes.rtcamp_like(statue_subdiv=3, detail=0.06, animated=8)
    gpu.upload(s)
    cam = s.camera
    for frame in (0, 3, 7):
        t = frame / 24.0
        xf, inv = s.transforms(t)
        fp = hj.frame_params(s, 96, 54, 2, frame=frame, max_depth=6, debug_paths=1)
        gpu.build_accel(xf, inv, mode=abi.ACCEL_BUILD)
        gpu.render(fp)
        ids_b = gpu.readback(abi.AOV_PATH_IDS).copy(); rad_b = gpu.readback(abi.AOV_PATH_RADIANCE).copy()
        if frame:
            # topology of the previous frame's build, boxes refitted to this frame's transforms
            gpu.build_accel(*s.transforms((frame - 1) / 24.0), mode=abi.ACCEL_BUILD)
            gpu.build_accel(xf, inv, mode=abi.ACCEL_REFIT)
            gpu.render(fp)
            assert np.array_equal(gpu.readback(abi.AOV_PATH_IDS), ids_b)            # traversal order must not matter
            assert np.array_equal(gpu.readback(abi.AOV_PATH_RADIANCE), rad_b)
        want = Oracle(s, xforms=xf, inv_xforms=inv).render(fp, debug=True)
        pt.compare_paths("C4 stand-in 96x54x2 depth 6 frame %d" % frame, rad_b, ids_b, want, exact=True)
    moved = np.abs(s.transforms(7 / 24.0)[0] - s.transforms(0.0)[0]).max()
    assert moved > 0.05


def test_c5_many_triangles_and_builders(gpu):
    s = scenes.icosphere_field(n_spheres=300, subdiv=3)       # ~390k triangles, 305 instances
    assert s.n_triangles > 350000
    rays = box_rays(s, 60000, 5)
    o = Oracle(s, use_bvh=True)
    p0, t0 = o.trace(rays)
    a0, _ = o.trace(rays, any_hit=True)
    gpu.upload(s)
    for builder in (0, 1):
        gpu.set_option("builder", builder)
        gpu.build_accel()
        p, t = gpu.trace(rays)
        assert np.array_equal(p, p0) and np.array_equal(t, t0), builder
        a, _ = gpu.trace(rays, any_hit=True)
        assert np.array_equal(a, a0)
        for leaf_max in (1, 2, 3):
            gpu.set_option("leaf_max", leaf_max); gpu.build_accel()
            p, t = gpu.trace(rays[:20000])
            assert np.array_equal(p, p0[:20000]) and np.array_equal(t, t0[:20000])
        gpu.set_option("leaf_max", 3)
        # leaf_split off: multi-triangle leaves keep one box even when the wide node has free slots
        gpu.set_option("leaf_split", 0); gpu.build_accel()
        p, t = gpu.trace(rays[:20000])
        assert np.array_equal(p, p0[:20000]) and np.array_equal(t, t0[:20000])
        gpu.set_option("leaf_split", 1)
        # PLOC with every round as separate launches (no single-block tail): same tree, same answers
        gpu.set_option("ploc_tail", 0); gpu.build_accel()
        p, t = gpu.trace(rays[:20000])
        assert np.array_equal(p, p0[:20000]) and np.array_equal(t, t0[:20000])
        gpu.set_option("ploc_tail", 1)
    gpu.set_option("builder", 0)
    assert (p0 != 0xffffffff).mean() > 0.5


def test_furnace_energy_conservation(gpu):
    """White-furnace intent of the reference's default config (render_option.json:11-12): a metallic sphere under a
    constant sky must not gain energy (the multiple-scattering GGX walk returns weights <= 1 for F0 = 1)."""
    for rough in (0.2, 0.6, 1.0):
        s = scenes.furnace_sphere(metallic=1.0, roughness=rough, subdiv=4)
        gpu.upload(s); gpu.build_accel()
        fp = hj.frame_params(s, 64, 64, 16, frame=0, max_depth=10, integrator=abi.INTEGRATOR_PT)
        gpu.render(fp)
        col = gpu.readback(abi.AOV_COLOR)[..., :3]
        centre = col[24:40, 24:40].mean()
        assert 0.85 < centre <= 1.0 + 1e-3, (rough, centre)
        want = Oracle(s).render(fp)
        assert pt.bit_equal(col, want["color"][..., :3])


@pytest.mark.parametrize("which", ["cornell", "c2", "c3", "no_lights"])
def test_mis_emitter_first_query_is_the_closest_hit_query(gpu, which):
    """The MIS emitter ray (rt.h:383-420) uses its closest hit only if that is a light. The emitter-first query (closest
    emitter, then any hit in front of it — the default outside debug_paths) must give bit-identical frames to the plain
    closest-hit query, whose per-path radiance the tests above pin against the oracle."""
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if which == "cornell":
        s = hj.load_npz(os.path.join(root, "tests", "golden", "cornelbox.npz")); s.set_sky((0.8, 0.8, 0.8)); depth = 8
    elif which == "c2":
        s = scenes.rtcamp_like(statue_subdiv=4, detail=0.08); depth = 10
    elif which == "c3":
        s = scenes.metaglass_room(subdiv=3); depth = 32
    else:                                  # no emitter at all: the emitter phase ends at the root, sky hits still count
        s = hj.load_npz(os.path.join(root, "tests", "golden", "cornelbox.npz")); s.set_sky((0.8, 0.8, 0.8)); depth = 6
        for m in s.materials:
            m["emission"] = (0.0, 0.0, 0.0); m["is_light"] = 0
        s.finalize_lights()
        assert s.light_prim_ids.size == 0
    fp = hj.frame_params(s, 160, 90, 4, frame=2, max_depth=depth, integrator=abi.INTEGRATOR_MIS)
    gpu.upload(s); gpu.build_accel()
    out = []
    for efirst in (0, 1):
        gpu.set_option("mis_emitter_first", efirst)
        gpu.render(fp)
        out.append(gpu.readback(abi.AOV_ACCUM).copy())
    gpu.set_option("mis_emitter_first", 1)
    assert np.array_equal(out[0].view(np.uint32), out[1].view(np.uint32))
    assert np.nansum(out[0][..., :3]) > 0
    # and after a refit (the emitter marks are rebuilt with the boxes)
    gpu.build_accel(mode=abi.ACCEL_REFIT)
    gpu.render(fp)
    assert np.array_equal(gpu.readback(abi.AOV_ACCUM).view(np.uint32), out[0].view(np.uint32))


@pytest.mark.parametrize("integrator", [abi.INTEGRATOR_PT, abi.INTEGRATOR_NEE, abi.INTEGRATOR_MIS])
def test_equirect_sky_image(gpu, integrator):
    """The miss program's equirect lookup (u = atan2(z,x)/2pi, v = acos(y)/pi, bilinear, wrap; setSky renderer.h:802-851)
    with a NON-constant sky image — round 1 only ever rendered 1x1 skies. A seeded 64x32 image with a 40x 'sun', intensity 0.75:
    a rough metal sphere and a diffuse floor under it (every miss goes through atan2f / acosf of hj_libm), and the Cornell box,
    whose open side lets rays escape. Per path the same bits as the oracle."""
    from henjou_renderer_b200 import hdrio
    sky = hdrio.rgbe_to_float4(hdrio.float_to_rgbe(hdrio.procedural_sky(64, 32, seed=4)))      # what the .hdr loader would deliver
    assert sky.max() > 20 and sky[..., :3].min() >= 0
    s = scenes.furnace_sphere(metallic=1.0, roughness=0.35, subdiv=3)
    s.set_sky_image(sky)
    fp = hj.frame_params(s, 96, 96, 3, frame=2, max_depth=8, integrator=integrator, ibl_intensity=0.75, debug_paths=1)
    gpu.upload(s); gpu.build_accel(); gpu.render(fp)
    want = Oracle(s).render(fp, debug=True)
    pt.compare_paths("equirect sky 64x32, metal sphere, integrator %d" % integrator, gpu.readback(abi.AOV_PATH_RADIANCE), gpu.readback(abi.AOV_PATH_IDS), want, exact=True)
    assert pt.bit_equal(gpu.readback(abi.AOV_COLOR), want["color"])
    miss = want["path_ids"][..., 0] == 0xffffffff
    assert miss.mean() > 0.25 and np.unique(want["path_radiance"][miss][:, 0]).size > 50      # primary misses see many different texels
    import os
    c = hj.load_npz(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "cornelbox.npz"))
    c.set_sky_image(sky)
    xf, inv = c.transforms(1 / 24)
    fp = hj.frame_params(c, 80, 80, 2, frame=1, max_depth=8, integrator=integrator, ibl_intensity=0.75, debug_paths=1)
    gpu.upload(c); gpu.build_accel(xf, inv); gpu.render(fp)
    want = Oracle(c, xforms=xf, inv_xforms=inv).render(fp, debug=True)
    pt.compare_paths("equirect sky 64x32, cornell box, integrator %d" % integrator, gpu.readback(abi.AOV_PATH_RADIANCE), gpu.readback(abi.AOV_PATH_IDS), want, exact=True)


def test_nonfinite_vertex_is_an_error_not_a_hang(gpu):
    """one NaN / inf vertex makes a cluster that can never merge (every merged area is NaN). Round 1's single-block PLOC
    tail then spun for ever. Now k_flatten flags non-finite world positions, both PLOC kernels stop when a round merges nothing, and
    hj_accel_build reports it."""
    import copy, os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    good = hj.load_npz(os.path.join(root, "tests", "golden", "cornelbox.npz"))
    for bad_value in (np.nan, np.inf):
        s = copy.copy(good)
        s.vertices = good.vertices.copy(); s.vertices[100, 1] = bad_value
        gpu.upload(s)
        for builder in (1, 0):
            gpu.set_option("builder", builder)
            with pytest.raises(hj.HenjouError, match="not finite"):
                gpu.build_accel()
        gpu.set_option("builder", 1)
    # a NaN in the instance matrices too
    gpu.upload(good)
    xf, inv = good.transforms(0.0)
    xf = xf.copy(); xf[1, 0, 3] = np.nan
    with pytest.raises(hj.HenjouError, match="not finite"):
        gpu.build_accel(xf, inv)
    gpu.build_accel(*good.transforms(0.0))                     # and the context is still usable
    gpu.render(hj.frame_params(good, 32, 32, 1))
    assert np.isfinite(gpu.readback(abi.AOV_COLOR)).all()


def test_async_accel_update_under_a_frame_in_flight(gpu):
    """hj_accel_build_async builds frame N+1's structure into the second buffer set on the build stream while frame N
    (enqueued with hj_render_async) is still being shaded; the next render switches to it. Every frame must be bit-identical to the
    blocking sequence (build -> render -> read back), for rebuild and for refit."""
    s = scenes.rtcamp_like(statue_subdiv=4, detail=0.08, animated=8)
    gpu.upload(s)
    frames = list(range(0, 6))
    fps = [hj.frame_params(s, 160, 90, 2, frame=f, max_depth=6, debug_paths=1) for f in frames]
    want = []
    for f, fp in zip(frames, fps):                             # blocking reference sequence
        gpu.build_accel(*s.transforms(f / 24.0), mode=abi.ACCEL_BUILD)
        gpu.render(fp)
        want.append([gpu.readback(a).copy() for a in (abi.AOV_ACCUM, abi.AOV_PATH_RADIANCE, abi.AOV_PATH_IDS, abi.AOV_RGBA8)])
    for mode in (abi.ACCEL_BUILD, abi.ACCEL_REFIT):
        gpu.build_accel(*s.transforms(frames[0] / 24.0), mode=abi.ACCEL_BUILD)
        for k, (f, fp) in enumerate(zip(frames, fps)):
            gpu.render(fp, sync=False)                         # frame f in flight ...
            if k + 1 < len(frames):
                gpu.build_accel(*s.transforms(frames[k + 1] / 24.0), mode=mode, sync=False)   # ... while f+1's structure is built
            gpu.sync()
            got = [gpu.readback(a).copy() for a in (abi.AOV_ACCUM, abi.AOV_PATH_RADIANCE, abi.AOV_PATH_IDS, abi.AOV_RGBA8)]
            for a, b in zip(got, want[k]):
                assert pt.bit_equal(a, b), (mode, f)
    # accel_wait reports a failed asynchronous build; the previous structure stays in place
    xf, inv = s.transforms(0.0)
    bad = xf.copy(); bad[0, 0, 0] = np.inf
    gpu.build_accel(bad, inv, sync=False)
    with pytest.raises(hj.HenjouError, match="not finite"):
        gpu.accel_wait()
    gpu.render(fps[-1])
    assert pt.bit_equal(gpu.readback(abi.AOV_ACCUM), want[-1][0])
    assert gpu.stats()["ms_accel"] > 0


def test_sky_importance_sampling_extension(gpu):
    """hj_frame_params::sky_sampling = 1 (MIS integrator) adds a light-sampled sky term with MIS — a luminance x sin(theta)
    table built on the GPU at upload, one more RNG dimension and one more shadow ray per non-specular vertex. Off by default (the
    reference BSDF-samples the sky only). (a) per path the same bits as the oracle's restatement of it; (b) it estimates the same
    image as the reference behaviour (two independent converged renders agree) with far less noise at equal sample counts."""
    from henjou_renderer_b200 import hdrio
    sky = hdrio.rgbe_to_float4(hdrio.float_to_rgbe(hdrio.procedural_sky(64, 32, seed=4, sun=400.0)))
    s = scenes.furnace_sphere(metallic=0.0, roughness=0.8, subdiv=3, basecolor=(0.8, 0.8, 0.8))
    s.set_sky_image(sky)
    gpu.upload(s); gpu.build_accel()
    fp = hj.frame_params(s, 96, 96, 3, frame=2, max_depth=6, integrator=abi.INTEGRATOR_MIS, ibl_intensity=0.75, sky_sampling=1, debug_paths=1)
    gpu.render(fp)
    want = Oracle(s).render(fp, debug=True)
    pt.compare_paths("sky_sampling=1, sun sky 64x32, diffuse sphere", gpu.readback(abi.AOV_PATH_RADIANCE), gpu.readback(abi.AOV_PATH_IDS), want, exact=True)
    st = gpu.stats()
    assert int(st["rays_shadow"]) == int(want["traced_shadow"]) and st["rays_shadow"] > 0
    # the rtcamp-like stand-in under the same sky: emissive triangles AND sky samples, thin-film / EC-GGX / glass classes
    c2 = scenes.rtcamp_like(statue_subdiv=3, detail=0.06)
    c2.set_sky_image(sky)
    gpu.upload(c2); gpu.build_accel()
    fp2 = hj.frame_params(c2, 96, 54, 2, frame=1, max_depth=10, integrator=abi.INTEGRATOR_MIS, sky_sampling=1, debug_paths=1)
    gpu.render(fp2)
    pt.compare_paths("sky_sampling=1, C2 stand-in", gpu.readback(abi.AOV_PATH_RADIANCE), gpu.readback(abi.AOV_PATH_IDS), Oracle(c2).render(fp2, debug=True), exact=True)
    # (b) unbiased against the reference behaviour, and the point of it: variance
    gpu.upload(s); gpu.build_accel()
    def img(spp, on, frame):
        gpu.render(hj.frame_params(s, 64, 64, spp, frame=frame, max_depth=6, integrator=abi.INTEGRATOR_MIS, sky_sampling=on))
        return gpu.readback(abi.AOV_COLOR)[..., :3].astype(np.float64)
    ref_on, ref_off = img(16384, 1, 1), img(16384, 0, 2)
    # about one path in 1e8 ends in NaN (a 0/0 inside the reference's BSDF code, reproduced on purpose): leave such pixels out
    ok = np.isfinite(ref_on).all(-1) & np.isfinite(ref_off).all(-1)
    assert ok.mean() > 0.999
    ref_on, ref_off = np.where(ok[..., None], ref_on, 0.0), np.where(ok[..., None], ref_off, 0.0)
    rel = lambda a, b: float(np.mean((((np.where(ok[..., None], a, 0.0) - b) ** 2).sum(-1) / ((ref_on ** 2).sum(-1) + 1e-2))[ok]))
    assert abs(ref_on.mean() - ref_off.mean()) <= 5e-3 * ref_on.mean(), (ref_on.mean(), ref_off.mean())
    assert rel(ref_off, ref_on) < 5e-3
    e_on, e_off = rel(img(16, 1, 3), ref_on), rel(img(16, 0, 3), ref_on)
    pt.log_measurement(case="sky_sampling: relMSE at 16 spp with / without (sun sky, diffuse sphere)", on=e_on, off=e_off, ratio=e_off / e_on,
                       converged_means=(ref_on.mean(), ref_off.mean()), converged_rel_mse=rel(ref_off, ref_on))
    assert e_on * 20 < e_off
    # a constant 1x1 sky is a valid table too (uniform in (u, v)); a black sky has no table and draws nothing but the dimension
    s.set_sky((0.0, 0.0, 0.0))
    gpu.upload(s); gpu.build_accel()
    fpb = hj.frame_params(s, 32, 32, 2, max_depth=4, integrator=abi.INTEGRATOR_MIS, sky_sampling=1, debug_paths=1)
    gpu.render(fpb)
    pt.compare_paths("sky_sampling=1, black 1x1 sky", gpu.readback(abi.AOV_PATH_RADIANCE), gpu.readback(abi.AOV_PATH_IDS), Oracle(s).render(fpb, debug=True), exact=True)


def test_nan_repair_option(gpu):
    """Option nan_repair (off by default: parity): non-finite samples — the reference's BSDF code yields about one per 1e8 paths, here
    forced by a NaN texel in the sky image — are left out of their pixel's mean instead of poisoning it. Pixels without such a sample
    are the same bits either way."""
    from henjou_renderer_b200 import hdrio
    sky = hdrio.rgbe_to_float4(hdrio.float_to_rgbe(hdrio.procedural_sky(64, 32, seed=4)))
    sky[10, 20, :3] = np.nan
    s = scenes.furnace_sphere(metallic=0.0, roughness=0.8, subdiv=3, basecolor=(0.8, 0.8, 0.8))
    s.set_sky_image(sky)
    gpu.upload(s); gpu.build_accel()
    fp = hj.frame_params(s, 96, 96, 8, frame=1, max_depth=5, integrator=abi.INTEGRATOR_MIS)
    gpu.render(fp)
    plain = gpu.readback(abi.AOV_COLOR).copy(); acc0 = gpu.readback(abi.AOV_ACCUM).copy()
    bad = ~np.isfinite(plain[..., :3]).all(-1)
    assert 0 < bad.mean() < 0.5 and np.all(acc0[..., 3] == 8)
    gpu.set_option("nan_repair", 1)
    try:
        gpu.render(fp)
        fixed = gpu.readback(abi.AOV_COLOR).copy(); acc1 = gpu.readback(abi.AOV_ACCUM).copy(); rgba = gpu.readback(abi.AOV_RGBA8).copy()
    finally:
        gpu.set_option("nan_repair", 0)
    assert np.isfinite(fixed).all()
    assert np.array_equal(fixed[~bad].view(np.uint32), plain[~bad].view(np.uint32))        # untouched where nothing was dropped
    assert np.all(acc1[..., 3][~bad] == 8) and np.all(acc1[..., 3][bad] < 8)                 # the count of samples kept
    assert rgba[bad][:, :3].max() > 0
