"""Per-path radiance parity at BASELINE.json's FULL configurations (round-1 verdict, item 1): the GPU renders the whole frame
at the stated size with debug_paths, the CPU oracle renders a band of pixel rows of the same frame (Oracle.render(rows=...)),
and the two are compared path by path:
  * against oracle/libhjo.so (restatement, same libm as the kernels): EXACT — the same bits;
  * against oracle/_ref (the reference's own headers on glibc's libm; max depth 10 only, rt.h:89): first-hit ids equal and the
    fraction of paths beyond 1e-4 relative within 2x of what the libm difference alone produces on the CPU
    (profiles/r02_parity_cpu.md; measured fractions of these tests: profiles/r02_parity.md).
C1 is rendered whole at its stated 512x512 / 16 spp / depth 8, with image-level numbers (RMSE, relMSE) and a converged-image
check between independent sample sets."""
import os

import numpy as np
import pytest

import henjou_renderer_b200 as hj
from henjou_renderer_b200 import abi, scenes
import oracle_binding as ob
from oracle_binding import Oracle
import parity_tools as pt

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def gpu():
    r = hj.Renderer(0)
    yield r
    r.close()


def _band(a, y0, y1):
    return {k: (v[:, y0:y1] if k.startswith("path_") else v[y0:y1]) for k, v in a.items() if isinstance(v, np.ndarray)}


def _gpu_paths(gpu, fp, y0, y1):
    gpu.render(fp)
    return gpu.readback(abi.AOV_PATH_RADIANCE)[:, y0:y1].copy(), gpu.readback(abi.AOV_PATH_IDS)[:, y0:y1].copy()


def _rel_mse(a, b):
    a, b = a[..., :3].astype(np.float64), b[..., :3].astype(np.float64)
    ok = np.isfinite(a).all(-1) & np.isfinite(b).all(-1)
    return float(np.mean(((a - b)[ok] ** 2) / (b[ok] ** 2 + 1e-2)))


def test_c1_whole_frame_512x512_16spp_depth8(gpu):
    s = hj.load_npz(os.path.join(ROOT, "tests", "golden", "cornelbox.npz")); s.set_sky((0.8, 0.8, 0.8))
    xf, inv = s.transforms(1 / 24)
    fp = hj.frame_params(s, 512, 512, 16, frame=1, max_depth=8, integrator=abi.INTEGRATOR_MIS, debug_paths=1)
    gpu.upload(s); gpu.build_accel(xf, inv)
    rad, ids = _gpu_paths(gpu, fp, 0, 512)
    want = Oracle(s, "restatement", xforms=xf, inv_xforms=inv).render(fp, debug=True)
    pt.compare_paths("C1 cornelbox 512x512x16 MIS depth 8 (whole frame) vs restatement", rad, ids, want, exact=True)
    col = gpu.readback(abi.AOV_COLOR)
    assert pt.bit_equal(col, want["color"]) and np.array_equal(gpu.readback(abi.AOV_RGBA8), want["rgba8"])
    pt.log_measurement(case="C1 image vs restatement", rmse=float(np.sqrt(np.mean((col[..., :3] - want["color"][..., :3]) ** 2))), rel_mse=_rel_mse(col, want["color"]))
    # the same frame at the reference's hard-wired depth 10, 4 spp, against the reference's own headers
    if ob.lib("reference") is not None:
        fp10 = hj.frame_params(s, 512, 512, 4, frame=1, max_depth=10, integrator=abi.INTEGRATOR_MIS, debug_paths=1)
        rad, ids = _gpu_paths(gpu, fp10, 0, 512)
        ref = Oracle(s, "reference", xforms=xf, inv_xforms=inv).render(fp10, debug=True)
        pt.compare_paths("C1 cornelbox 512x512x4 MIS depth 10 (whole frame) vs reference headers", rad, ids, ref, exact=False, bound=1e-4)
        col = gpu.readback(abi.AOV_COLOR)
        rmse = float(np.sqrt(np.mean((col[..., :3] - ref["color"][..., :3]) ** 2)))
        pt.log_measurement(case="C1 image vs reference headers (same seeds)", rmse=rmse, rel_mse=_rel_mse(col, ref["color"]),
                           rgba8_differing=float((gpu.readback(abi.AOV_RGBA8) != ref["rgba8"]).any(-1).mean()))
        assert rmse <= 2e-3, rmse                   # RMSE <= 2e-3 linear, same seeds


def test_c1_converged_image_independent_seeds(gpu):
    """north_star: 'converged images must reach RMSE/FLIP below a stated tolerance'. Two INDEPENDENT sample sets (different CMJ
    scramble = frame number): 4096 spp on the GPU against 256 spp of the CPU oracle at 128x128. Stated tolerance: the relative MSE
    between the two must be within 1.5x of the oracle's own 256-spp noise (estimated from two of its 128-spp halves / 2... i.e.
    half-set difference / 4) and the mean luminance within 0.5 % (CPU runs with other scrambles and sample counts scatter by 0.15 %):
    a biased estimator fails both however many samples it takes."""
    s = hj.load_npz(os.path.join(ROOT, "tests", "golden", "cornelbox.npz")); s.set_sky((0.8, 0.8, 0.8))
    xf, inv = s.transforms(1 / 24)
    gpu.upload(s); gpu.build_accel(xf, inv)
    W = H = 128
    gpu.render(hj.frame_params(s, W, H, 4096, frame=11, max_depth=8, integrator=abi.INTEGRATOR_MIS))
    big = gpu.readback(abi.AOV_COLOR)[..., :3].astype(np.float64)
    o = Oracle(s, "restatement", xforms=xf, inv_xforms=inv)
    h1 = o.render(hj.frame_params(s, W, H, 256, frame=5, max_depth=8, integrator=abi.INTEGRATOR_MIS, sample_begin=0, sample_end=128))["accum"][..., :3].astype(np.float64) / 128
    h2 = o.render(hj.frame_params(s, W, H, 256, frame=5, max_depth=8, integrator=abi.INTEGRATOR_MIS, sample_begin=128, sample_end=256))["accum"][..., :3].astype(np.float64) / 128
    cpu = 0.5 * (h1 + h2)
    # robust to fireflies: relative squared error with the converged image in the denominator, 99th-percentile trimmed
    def rel(a, b):
        e = ((a - b) ** 2).sum(-1) / ((big ** 2).sum(-1) + 1e-2)
        return float(np.mean(np.sort(e.ravel())[: int(0.99 * e.size)]))
    noise_256 = rel(h1, h2) / 4.0                    # Var(mean of halves) = Var(half difference) / 4
    got = rel(cpu, big)
    lum = lambda im: float(np.mean(im @ np.array([0.2126, 0.7152, 0.0722])))   # the mean is the unbiased statistic (the median of a noisy image moves with spp)
    pt.log_measurement(case="C1 converged: GPU 4096 spp (scramble 11) vs CPU 256 spp (scramble 5), 128x128", rel_mse=got, cpu_noise_rel_mse=noise_256,
                       ratio=got / noise_256, mean_luminance=(lum(big), lum(cpu)))
    assert got <= 1.5 * noise_256 + 1e-5, (got, noise_256)
    assert abs(lum(big) - lum(cpu)) <= 5e-3 * lum(cpu)


@pytest.fixture(scope="module")
def c2():
    return scenes.rtcamp_like()


def test_c2_1080p_band_per_path(gpu, c2):
    """The real C2: 1.85 M triangles, 1920x1080, MIS, depth 10, thin-film on; 32 rows x 4 spp = 245 760 paths."""
    assert c2.n_triangles > 1_800_000
    xf, inv = c2.transforms(1 / 24.0)
    fp = hj.frame_params(c2, 1920, 1080, 4, frame=1, max_depth=10, integrator=abi.INTEGRATOR_MIS, debug_paths=1)
    gpu.upload(c2); gpu.build_accel(xf, inv)
    gpu.render(fp)
    grad, gids = gpu.readback(abi.AOV_PATH_RADIANCE), gpu.readback(abi.AOV_PATH_IDS)
    o = Oracle(c2, "restatement", xforms=xf, inv_xforms=inv)
    oref = Oracle(c2, "reference", xforms=xf, inv_xforms=inv) if ob.lib("reference") is not None else None
    for y0 in (380, 700):                            # statue + temple columns / floor + emissive ring reflections
        y1 = y0 + 16
        rad, ids = grad[:, y0:y1], gids[:, y0:y1]
        want = _band(o.render(fp, rows=(y0, y1), debug=True), y0, y1)
        pt.compare_paths("C2 1920x1080x4 rows %d..%d vs restatement" % (y0, y1), rad, ids, want, exact=True)
        if oref is not None:
            ref = _band(oref.render(fp, rows=(y0, y1), debug=True), y0, y1)
            pt.compare_paths("C2 1920x1080x4 rows %d..%d vs reference headers" % (y0, y1), rad, ids, ref, exact=False, bound=2e-3)
            if y0 == 380:                               # the bridge at full size: the glibc build of the restatement IS the reference, bit for bit
                gl = _band(Oracle(c2, "glibc", xforms=xf, inv_xforms=inv).render(fp, rows=(y0, y1), debug=True), y0, y1)
                assert pt.bit_equal(gl["path_radiance"][..., :3], ref["path_radiance"][..., :3])


def test_c3_1080p_depth32_band_per_path(gpu):
    s = scenes.metaglass_room()
    fp = hj.frame_params(s, 1920, 1080, 2, frame=1, max_depth=32, integrator=abi.INTEGRATOR_MIS, debug_paths=1)
    gpu.upload(s); gpu.build_accel()
    y0, y1 = 524, 556
    rad, ids = _gpu_paths(gpu, fp, y0, y1)
    want = _band(Oracle(s, "restatement").render(fp, rows=(y0, y1), debug=True), y0, y1)
    pt.compare_paths("C3 1920x1080x2 depth 32 rows %d..%d vs restatement" % (y0, y1), rad, ids, want, exact=True)
    assert (want["path_radiance"][..., 3] > 20).any()
    if ob.lib("reference") is not None:              # the reference's integrators stop at depth 10
        fp10 = hj.frame_params(s, 1920, 1080, 2, frame=1, max_depth=10, integrator=abi.INTEGRATOR_MIS, debug_paths=1)
        rad, ids = _gpu_paths(gpu, fp10, y0, y1)
        ref = _band(Oracle(s, "reference").render(fp10, rows=(y0, y1), debug=True), y0, y1)
        pt.compare_paths("C3 1920x1080x2 depth 10 rows %d..%d vs reference headers" % (y0, y1), rad, ids, ref, exact=False, bound=2e-4)


def test_c5_4k_band_per_path(gpu):
    s = scenes.icosphere_field()
    assert s.n_triangles > 10_000_000
    fp = hj.frame_params(s, 3840, 2160, 2, frame=1, max_depth=10, integrator=abi.INTEGRATOR_MIS, debug_paths=1)
    gpu.upload(s); gpu.build_accel()
    y0, y1 = 1076, 1084
    rad, ids = _gpu_paths(gpu, fp, y0, y1)
    o = Oracle(s, "restatement")
    want = _band(o.render(fp, rows=(y0, y1), debug=True), y0, y1)
    pt.compare_paths("C5 3840x2160x2 rows %d..%d vs restatement" % (y0, y1), rad, ids, want, exact=True)
    o.close()
    if ob.lib("reference") is not None:
        ref = _band(Oracle(s, "reference").render(fp, rows=(y0, y1), debug=True), y0, y1)
        # the icosphere field is mostly rough metal (EC-GGX height walk, pow(1-U, 1/Lambda)): the most libm-sensitive of the scenes.
        # What the libm difference alone does to this band is measured on the CPU in the same breath (restatement on glibc vs on
        # hj_libm); since GPU == restatement and glibc build == reference, bit for bit, the two fractions are the same number.
        pt.compare_paths("C5 3840x2160x2 rows %d..%d vs reference headers" % (y0, y1), rad, ids, ref, exact=False, bound=2e-2)
        gl = _band(Oracle(s, "glibc").render(fp, rows=(y0, y1), debug=True), y0, y1)
        assert pt.bit_equal(gl["path_radiance"][..., :3], ref["path_radiance"][..., :3])       # the bridge holds at full size too
