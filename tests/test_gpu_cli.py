"""The headless C++ host end to end on the GPU: render_option.json + glTF in, `<image_name>_<NNN>.png` out
(reference: Renderer::initializeAndRender, renderer.h:1053-1318), against the Python driver rendering the same fixture
through the same C ABI. The two hosts differ only in last-ulp host arithmetic (float cofactor inverse vs numpy's double
inverse; tanf vs tan), so the images must agree except for a handful of 8-bit roundings."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import henjou_renderer_b200 as hj                      # noqa: E402
from henjou_renderer_b200 import abi                   # noqa: E402
import npz_to_gltf                                     # noqa: E402

CLI = os.path.join(ROOT, "henjou-renderer_b200", "henjou_b200_cli")
pytestmark = pytest.mark.gpu


def python_frames(scene, W, H, spp, frames, fps, **kw):
    r = hj.Renderer(0)
    r.upload(scene)
    out = []
    for i, fr in enumerate(frames):
        t = float(np.float32(fr) / np.float32(fps))
        xf, inv = scene.transforms(t)
        r.build_accel(xf, inv, mode=abi.ACCEL_BUILD if i == 0 else abi.ACCEL_REFIT)
        r.render(hj.frame_params(scene, W, H, spp, frame=fr, fps=fps, **kw))
        out.append(r.readback(abi.AOV_RGBA8)[::-1].copy())      # the PNG is written bottom-up (sutil::saveImage)
    r.close()
    return out


@pytest.mark.parametrize("fixture", ["cornelbox", "cornelbox_texture_test"])
def test_cli_renders_the_same_png(tmp_path, fixture):
    scene = hj.load_npz(os.path.join(ROOT, "tests", "golden", fixture + ".npz"))
    scene.set_sky((0.8, 0.8, 0.8))
    W, H, spp, frames = 160, 120, 8, [1, 2]
    npz_to_gltf.export_gltf(scene, str(tmp_path / "Model"), fixture)
    cam = scene.camera
    kw = {}
    if cam is None:         # no camera node in the file: the JSON camera is used
        cam = dict(position=np.float32([5.4, 0.0, 0.0]), direction=np.float32([-1, 0, 0]), fov=float(np.float32(np.pi * 30.0 / 180.0)), animation=None)
        kw = dict(camera_position=(5.4, 0.0, 0.0), camera_direction=(-1, 0, 0), camera_fov=30.0)
    npz_to_gltf.write_render_option(str(tmp_path / "render_option.json"), "./Model/", fixture + ".gltf", "img", W, H, spp, fps=24,
                                    start_frame=frames[0], end_frame=frames[-1] + 1, sky=(0.8, 0.8, 0.8), **kw)
    r = subprocess.run([CLI, "render_option.json"], cwd=str(tmp_path), capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    want = python_frames(scene, W, H, spp, frames, 24, camera=cam)
    for fr, w in zip(frames, want):
        got = npz_to_gltf.read_png(str(tmp_path / ("img_%03d.png" % fr)))
        assert got.shape == w.shape
        diff = np.abs(got.astype(np.int32) - w.astype(np.int32))
        # a last-ulp difference in a camera ray can flip a discrete decision (lobe choice, roulette) of a single sample:
        # few pixels may differ, by any amount; the image as a whole may not
        assert (diff > 0).any(-1).mean() < 5e-3 and diff.mean() < 0.02, "frame %d: %g of the pixels differ, mean |d| %g" % (fr, (diff > 0).any(-1).mean(), diff.mean())
        assert w[..., :3].mean() > 5         # not a black frame


def test_cli_without_gpu_library_fails_loudly(tmp_path):
    """A scene without triangles / a missing scene file must end with a message and a non-zero code, never a blank PNG."""
    npz_to_gltf.write_render_option(str(tmp_path / "render_option.json"), "./Model/", "nothing.gltf", "img", 16, 16, 1)
    r = subprocess.run([CLI, "render_option.json"], cwd=str(tmp_path), capture_output=True, text=True, timeout=60)
    assert r.returncode != 0 and "cannot open" in r.stderr
    assert not list(tmp_path.glob("*.png"))


def test_cli_render_modes(tmp_path):
    """Render_mode (render_json_loader.h:115-136, renderer.h:1089-1115): Denoise = hj_post on the same frame;
    DenoiseUpScale2X renders at half the resolution and writes the full-size image."""
    scene = hj.load_npz(os.path.join(ROOT, "tests", "golden", "cornelbox.npz"))
    scene.set_sky((0.8, 0.8, 0.8))
    W, H, spp = 96, 64, 4
    npz_to_gltf.export_gltf(scene, str(tmp_path / "Model"), "c")
    r = hj.Renderer(0)
    r.upload(scene)
    xf, inv = scene.transforms(float(np.float32(1) / np.float32(24)))
    for mode, post, (w, h) in (("Denoise", abi.POST_DENOISE, (W, H)), ("DenoiseUpScale2X", abi.POST_DENOISE_UPSCALE2X, (W // 2, H // 2))):
        npz_to_gltf.write_render_option(str(tmp_path / "render_option.json"), "./Model/", "c.gltf", mode, W, H, spp, fps=24, start_frame=1,
                                        end_frame=2, render_mode=mode)
        p = subprocess.run([CLI, "render_option.json"], cwd=str(tmp_path), capture_output=True, text=True, timeout=300)
        assert p.returncode == 0, p.stderr
        got = npz_to_gltf.read_png(str(tmp_path / (mode + "_001.png")))
        assert got.shape == (H, W, 4)
        r.build_accel(xf, inv)
        r.render(hj.frame_params(scene, w, h, spp, frame=1, fps=24, camera=scene.camera))
        r.post(post)
        want = r.readback(abi.AOV_OUTPUT_RGBA8)[::-1]
        diff = np.abs(got.astype(np.int32) - want.astype(np.int32))
        assert (diff > 1).any(-1).mean() < 2e-2 and diff.mean() < 0.05, (mode, (diff > 1).any(-1).mean(), diff.mean())
    r.close()


def test_cli_time_limit_and_frame_sequence(tmp_path):
    """Frames are written by a worker while the next one renders: every PNG of the sequence must be there and differ
    (animated camera); --enforce-time-limit applies the check the reference carries commented out (renderer.h:1247-1254)."""
    scene = hj.load_npz(os.path.join(ROOT, "tests", "golden", "cornelbox.npz"))
    npz_to_gltf.export_gltf(scene, str(tmp_path / "Model"), "c")
    npz_to_gltf.write_render_option(str(tmp_path / "render_option.json"), "./Model/", "c.gltf", "seq", 64, 48, 2, fps=24, start_frame=1, end_frame=6)
    p = subprocess.run([CLI, "render_option.json"], cwd=str(tmp_path), capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    imgs = [npz_to_gltf.read_png(str(tmp_path / ("seq_%03d.png" % f))) for f in range(1, 6)]
    assert all(im.shape == (48, 64, 4) for im in imgs)
    assert all((imgs[i] != imgs[i + 1]).any() for i in range(4))          # frame index = RNG scramble: no two frames alike
    npz_to_gltf.write_render_option(str(tmp_path / "render_option.json"), "./Model/", "c.gltf", "lim", 64, 48, 2, fps=24, start_frame=1, end_frame=6,
                                    time_limit=0.0)
    p = subprocess.run([CLI, "render_option.json", "--enforce-time-limit"], cwd=str(tmp_path), capture_output=True, text=True, timeout=300)
    assert p.returncode == 0 and "Over Time Limit" in p.stdout + p.stderr
    assert sorted(f.name for f in tmp_path.glob("lim_*.png")) == ["lim_001.png"]
