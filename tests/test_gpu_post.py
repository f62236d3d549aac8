"""The post stage (hj_post: what OptixDenoiserManager::denoise + float4ConvertColor do per frame, renderer.h:1256-1283)
against its CPU restatement oracle/post_ref.py. Pass-through (Render_mode "Default") must be exact; the a-trous stand-in
for the OptiX network agrees to 1e-5 (expf is the only operation that may round differently) and must actually denoise:
a 4 spp frame moves closer to the 512 spp frame."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import henjou_renderer_b200 as hj                      # noqa: E402
from henjou_renderer_b200 import abi                   # noqa: E402
import post_ref                                        # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    r = hj.Renderer(0)
    yield r
    r.close()


@pytest.fixture(scope="module")
def scene():
    s = hj.load_npz(os.path.join(ROOT, "tests", "golden", "cornelbox_texture_test.npz"))
    s.set_sky((0.8, 0.8, 0.8))
    return s


CAM = dict(position=np.float32([5.4, 0.0, 0.0]), direction=np.float32([-1, 0, 0]), fov=float(np.float32(np.pi * 30.0 / 180.0)), animation=None)


def render(gpu, scene, W, H, spp):
    gpu.upload(scene); gpu.build_accel()
    gpu.render(hj.frame_params(scene, W, H, spp, frame=1, camera=scene.camera or CAM))
    return [gpu.readback(a).copy() for a in (abi.AOV_COLOR, abi.AOV_ALBEDO, abi.AOV_NORMAL)]


def test_passthrough_is_the_colour_aov(gpu, scene):
    color, _, _ = render(gpu, scene, 97, 61, 4)
    with pytest.raises(hj.HenjouError):
        gpu.readback(abi.AOV_OUTPUT)                       # hj_post not called for this frame
    gpu.post(abi.POST_PASSTHROUGH)
    assert np.array_equal(gpu.readback(abi.AOV_OUTPUT).view(np.uint32), color.view(np.uint32))
    assert np.array_equal(gpu.readback(abi.AOV_OUTPUT_RGBA8), gpu.readback(abi.AOV_RGBA8))
    st = gpu.stats()
    assert st["post_launches"] == 1 and st["ms_post"] > 0


@pytest.mark.parametrize("size", [(97, 61), (160, 120)])
def test_denoise_matches_cpu_restatement(gpu, scene, size):
    W, H = size
    color, albedo, normal = render(gpu, scene, W, H, 4)
    gpu.post(abi.POST_DENOISE)
    got = gpu.readback(abi.AOV_OUTPUT)
    want = post_ref.denoise(color, albedo, normal)
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6)
    assert gpu.stats()["post_launches"] == 6               # 5 a-trous passes + the RGBA8 conversion
    q = gpu.readback(abi.AOV_OUTPUT_RGBA8).astype(np.int32)
    assert np.abs(q - post_ref.to_rgba8(got).astype(np.int32)).max() <= 1     # powf may round differently on a bin edge
    assert (q != post_ref.to_rgba8(got)).mean() < 1e-3


def test_denoise_thinfilm_scene_with_sky(gpu):
    """The C2 stand-in (thin-film, metals, sky pixels with zero albedo/normal guides); non-finite samples, if the frame has
    any, must come out repaired the same way on both sides."""
    from henjou_renderer_b200 import scenes
    s = scenes.rtcamp_like(statue_subdiv=3, detail=0.1)
    gpu.upload(s); gpu.build_accel()
    gpu.render(hj.frame_params(s, 128, 72, 2, frame=1, max_depth=10))
    color, albedo, normal = [gpu.readback(a).copy() for a in (abi.AOV_COLOR, abi.AOV_ALBEDO, abi.AOV_NORMAL)]
    gpu.post(abi.POST_DENOISE)
    np.testing.assert_allclose(gpu.readback(abi.AOV_OUTPUT), post_ref.denoise(color, albedo, normal), rtol=1e-5, atol=1e-6)


def test_upscale2x(gpu, scene):
    W, H = 81, 45                                          # odd sizes: the clamped border taps are exercised
    color, albedo, normal = render(gpu, scene, W, H, 4)
    gpu.post(abi.POST_DENOISE_UPSCALE2X)
    got = gpu.readback(abi.AOV_OUTPUT)
    assert got.shape == (2 * H, 2 * W, 4)
    want = post_ref.upscale2x(post_ref.denoise(color, albedo, normal))
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6)
    assert gpu.readback(abi.AOV_OUTPUT_RGBA8).shape == (2 * H, 2 * W, 4)
    # options: zero iterations = plain enlargement of the colour AOV, bit-exact
    gpu.set_option("denoise_iterations", 0)
    gpu.post(abi.POST_DENOISE_UPSCALE2X)
    assert np.array_equal(gpu.readback(abi.AOV_OUTPUT).view(np.uint32), post_ref.upscale2x(color).view(np.uint32))
    gpu.set_option("denoise_iterations", 5)
    with pytest.raises(hj.HenjouError):
        gpu.set_option("denoise_iterations", 9)
    with pytest.raises(hj.HenjouError):
        gpu.post(7)


def test_denoise_reduces_error(gpu):
    """4 spp of the C2 stand-in (emissive ring, thin-film, metals: noisy) vs 256 spp, relative MSE (Rousselle et al.)."""
    from henjou_renderer_b200 import scenes
    s = scenes.rtcamp_like(statue_subdiv=3, detail=0.1)
    W, H = 192, 108
    gpu.upload(s); gpu.build_accel()
    gpu.render(hj.frame_params(s, W, H, 256, frame=1, max_depth=10))
    ref = gpu.readback(abi.AOV_COLOR).copy()
    gpu.render(hj.frame_params(s, W, H, 4, frame=1, max_depth=10))
    noisy = gpu.readback(abi.AOV_COLOR).copy()
    gpu.post(abi.POST_DENOISE)
    den = gpu.readback(abi.AOV_OUTPUT)
    good = np.isfinite(ref).all(-1) & np.isfinite(noisy).all(-1)          # 0/0 samples the reference's BSDFs share

    def rel_mse(a):
        d = a[..., :3][good] - ref[..., :3][good]
        return float(np.mean(d * d / (ref[..., :3][good] ** 2 + 1e-2)))
    assert np.isfinite(den).all()
    assert rel_mse(noisy) > 0.05 and rel_mse(den) < 0.3 * rel_mse(noisy), (rel_mse(den), rel_mse(noisy))
