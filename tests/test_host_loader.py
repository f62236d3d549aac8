"""Headless C++ host (henjou-renderer_b200/host): render_option.json + glTF ingest against the Python mirror of the
reference loader (tools/gltf_to_npz.py, scene.py, camera.py — each a restatement of gltfloader.h / animation.h /
renderer.h:1145-1169). CPU only: `henjou_b200_cli --dump-scene` stops before any CUDA call.

Bar: every SceneData array, material record, decoded texture and light list bit-exact; per-frame instance matrices
bit-exact; inverse rows within 1e-6 (glm's float cofactor order is not reproduced, see hj_host_scene.h); camera vectors
bit-exact and camera_f within 1 ulp (C++ tanf vs Python's double tan)."""
import ctypes
import os
import struct
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import henjou_renderer_b200 as hj                      # noqa: E402
from henjou_renderer_b200 import abi, camera          # noqa: E402
import npz_to_gltf                                     # noqa: E402

CLI = os.path.join(ROOT, "henjou-renderer_b200", "henjou_b200_cli")
REF_GLTF = "/root/reference/HenjouRenderer/Model/test_gltf"


def read_dump(path):
    d = open(path, "rb").read()
    assert d[:8] == b"HJSCENE1"
    out, p = {}, 8
    while p < len(d):
        name = d[p:p + 24].split(b"\0")[0].decode()
        dt = chr(d[p + 24])
        (n,) = struct.unpack("<Q", d[p + 25:p + 33])
        p += 33
        np_dt = {"f": np.float32, "u": np.uint32, "i": np.int32, "b": np.uint8}[dt]
        size = n * np.dtype(np_dt).itemsize
        out[name] = np.frombuffer(d[p:p + size], np_dt).copy()
        p += size
    return out


def run_cli(args, cwd):
    r = subprocess.run([CLI] + args, cwd=cwd, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    return r


def materials_of(dump):
    raw = dump["materials"]
    n = raw.size // ctypes.sizeof(abi.hj_material)
    return (abi.hj_material * n).from_buffer_copy(raw.tobytes())


def check_against(scene, dump, frames, fps):
    assert np.array_equal(dump["vertices"].view(np.uint32), scene.vertices.astype(np.float32).reshape(-1).view(np.uint32))
    assert np.array_equal(dump["normals"].view(np.uint32), scene.normals.astype(np.float32).reshape(-1).view(np.uint32))
    assert np.array_equal(dump["texcoords"].view(np.uint32), scene.texcoords.astype(np.float32).reshape(-1).view(np.uint32))
    for k in ("indices", "material_ids", "prim_offsets", "light_prim_ids"):
        assert np.array_equal(dump[k], np.asarray(getattr(scene, k), np.uint32).reshape(-1)), k
    assert np.array_equal(dump["light_prim_emission"], scene.light_prim_emission.astype(np.float32).reshape(-1))
    mats = materials_of(dump)
    assert len(mats) == len(scene.materials)
    for got, want in zip(mats, scene.materials):
        for k, _ in abi.hj_material._fields_:
            g = getattr(got, k)
            if k in ("basecolor", "emission", "specular"):
                assert [np.float32(x) for x in g] == [np.float32(x) for x in want[k]], k
            elif isinstance(g, float):
                assert np.float32(g) == np.float32(want[k]), k
            else:
                assert int(g) == int(want[k]), k
    for i, (img, srgb) in enumerate(scene.textures):
        w, h, s = dump["tex%d_whs" % i]
        assert (h, w) == img.shape[:2] and s == srgb
        assert np.array_equal(dump["tex%d" % i].reshape(h, w, 4), img)
    for fr in frames:
        t = float(np.float32(fr) / np.float32(fps))
        xf, inv = scene.transforms(t)
        assert np.array_equal(dump["xf_%d" % fr].view(np.uint32), xf.reshape(-1).view(np.uint32)), "instance matrices, frame %d" % fr
        assert np.allclose(dump["inv_%d" % fr], inv.reshape(-1), rtol=1e-6, atol=1e-6)
        cam = scene.camera
        if cam is None:                 # no camera node: the JSON camera is used (test_fps_file_and_static_camera)
            continue
        pos, d, up, right, cf = camera.camera_basis(cam["position"], cam["direction"], cam["fov"], cam.get("animation"), t)
        got = dump["cam_%d" % fr]
        assert np.array_equal(got[:12], np.concatenate([pos, d, up, right]).astype(np.float32)), "camera basis, frame %d" % fr
        assert abs(got[12] - cf) <= 2 * np.spacing(np.float32(cf))


@pytest.mark.parametrize("fixture", ["cornelbox", "cornelbox_texture_test"])
def test_gltf_roundtrip_matches_fixture(tmp_path, fixture):
    """npz fixture -> glTF (tools/npz_to_gltf.py) -> C++ loader == the fixture."""
    scene = hj.load_npz(os.path.join(ROOT, "tests", "golden", fixture + ".npz"))
    npz_to_gltf.export_gltf(scene, str(tmp_path / "Model"), fixture)
    npz_to_gltf.write_render_option(str(tmp_path / "render_option.json"), "./Model/", fixture + ".gltf", "out", 64, 48, 4, fps=24,
                                    start_frame=0, end_frame=3)
    run_cli(["render_option.json", "--dump-scene", "scene.bin", "--quiet"], str(tmp_path))
    dump = read_dump(str(tmp_path / "scene.bin"))
    check_against(scene, dump, range(0, 3), 24)
    opt = dump["options"]
    assert (opt[0], opt[1], opt[2], opt[3]) == (64, 48, 4, 24)
    assert np.array_equal(dump["ibl"], np.array([0.8, 0.8, 0.8, 0.0], np.float32))


@pytest.mark.skipif(not os.path.isdir(REF_GLTF), reason="needs the reference's glTF files (build container only)")
@pytest.mark.parametrize("name", ["cornelbox", "cornelbox_texture_test"])
def test_reference_gltf_files(tmp_path, name):
    """The reference's own scenes through the C++ loader == through the Python mirror of gltfloader.h (PNG decode included)."""
    import gltf_to_npz
    scene = gltf_to_npz.load_gltf(os.path.join(REF_GLTF, name + ".gltf"))
    npz_to_gltf.write_render_option(str(tmp_path / "render_option.json"), REF_GLTF + "/", name + ".gltf", "out", 32, 32, 1, fps=24,
                                    start_frame=1, end_frame=3)
    run_cli(["render_option.json", "--dump-scene", "scene.bin", "--quiet"], str(tmp_path))
    check_against(scene, read_dump(str(tmp_path / "scene.bin")), range(1, 3), 24)


def test_fps_file_and_static_camera(tmp_path):
    """./fps.txt overrides Animation.fps (render_json_loader.h:16-60,162-169); without a camera node or with
    allow_camera_animation off the JSON camera is used un-normalised (renderer.h:1163-1169)."""
    scene = hj.load_npz(os.path.join(ROOT, "tests", "golden", "cornelbox.npz"))
    npz_to_gltf.export_gltf(scene, str(tmp_path / "Model"), "c")
    npz_to_gltf.write_render_option(str(tmp_path / "render_option.json"), "./Model/", "c.gltf", "out", 32, 32, 1, fps=24, start_frame=2, end_frame=3,
                                    allow_camera_animation=False, camera_position=(1, 2, 3), camera_direction=(0, 0, 2), camera_fov=30.0)
    (tmp_path / "fps.txt").write_text("30")
    run_cli(["render_option.json", "--dump-scene", "scene.bin", "--quiet"], str(tmp_path))
    dump = read_dump(str(tmp_path / "scene.bin"))
    assert dump["options"][3] == 30 and dump["options"][5] == -1
    fov = np.float32(np.pi * 30.0 / 180.0)
    assert dump["options"][4] == fov
    cam = dump["cam_2"]
    assert list(cam[:6]) == [1, 2, 3, 0, 0, 2]
    right = np.cross(np.float32([0, 0, 2]), np.float32([0, 1, 0])).astype(np.float32)
    up = np.cross(right, np.float32([0, 0, 2])).astype(np.float32)
    assert np.array_equal(cam[6:9], up) and np.array_equal(cam[9:12], right)


def test_bad_inputs_fail_loudly(tmp_path):
    r = subprocess.run([CLI, "missing.json", "--dump-scene", "x.bin"], cwd=str(tmp_path), capture_output=True, text=True)
    assert r.returncode == 1 and "cannot open" in r.stderr
    (tmp_path / "render_option.json").write_text('{"Image": {"image_width": 4}}')
    r = subprocess.run([CLI, "render_option.json", "--dump-scene", "x.bin"], cwd=str(tmp_path), capture_output=True, text=True)
    assert r.returncode == 1 and "missing key" in r.stderr


@pytest.mark.parametrize("rle", [True, False])
def test_hdr_sky_through_the_cpp_host(tmp_path, rle):
    """Sky.use_IBL: a Radiance .hdr (new-style RLE and flat scanlines) written by the Python mirror, read by host/hj_hdr.h through the
    CLI: the float4 sky image must be rgb * 2^(e-136), w = 0 (stb_image's decode, texture.h:67-101), bit for bit."""
    from henjou_renderer_b200 import hdrio
    scene = hj.load_npz(os.path.join(ROOT, "tests", "golden", "cornelbox.npz"))
    npz_to_gltf.export_gltf(scene, str(tmp_path / "Model"), "c")
    sky = hdrio.procedural_sky(64, 32, seed=4)
    sky[5, 8:40] = sky[5, 8]                                  # a long constant run for the RLE coder
    os.makedirs(str(tmp_path / "HDRI"))
    rgbe = hdrio.write_hdr(str(tmp_path / "HDRI" / "sky.hdr"), sky, rle=rle)
    npz_to_gltf.write_render_option(str(tmp_path / "render_option.json"), "./Model/", "c.gltf", "out", 32, 32, 1, ibl_path="./HDRI/sky.hdr", ibl_intensity=0.75)
    run_cli(["render_option.json", "--dump-scene", "scene.bin", "--quiet"], str(tmp_path))
    dump = read_dump(str(tmp_path / "scene.bin"))
    want = hdrio.rgbe_to_float4(rgbe)
    assert dump["ibl"].size == 64 * 32 * 4
    assert np.array_equal(dump["ibl"].view(np.uint32), want.reshape(-1).view(np.uint32))
    assert (np.abs(want[..., :3] - sky) <= sky.max(-1, keepdims=True) / 128 + 1e-6).all() and (want[..., 3] == 0).all()   # RGBE: 8 bits of the largest component
    assert dump["options"][6] == np.float32(0.75)
    if rle:
        assert os.path.getsize(str(tmp_path / "HDRI" / "sky.hdr")) < 64 * 32 * 4 + 60             # the runs really are coded
    # a missing file falls back to the constant default sky (texture.h:91-101)
    npz_to_gltf.write_render_option(str(tmp_path / "render_option.json"), "./Model/", "c.gltf", "out", 32, 32, 1, ibl_path="./HDRI/absent.hdr", sky=(0.1, 0.2, 0.3))
    run_cli(["render_option.json", "--dump-scene", "scene.bin", "--quiet"], str(tmp_path))
    assert np.array_equal(read_dump(str(tmp_path / "scene.bin"))["ibl"], np.array([0.1, 0.2, 0.3, 0.0], np.float32))


def test_loader_validates_accessor_shapes(tmp_path):
    """a NORMAL / TEXCOORD_0 / POSITION accessor that is not FLOAT VECn, or animation keys shorter than their data, must be
    refused (they would be read past the validated range); a mesh node without triangles is dropped instead of breaking the scene."""
    import json
    scene = hj.load_npz(os.path.join(ROOT, "tests", "golden", "cornelbox.npz"))
    npz_to_gltf.export_gltf(scene, str(tmp_path / "Model"), "c")
    npz_to_gltf.write_render_option(str(tmp_path / "render_option.json"), "./Model/", "c.gltf", "out", 32, 32, 1)
    path = str(tmp_path / "Model" / "c.gltf")
    good = json.load(open(path))
    prim = good["meshes"][0]["primitives"][0]
    for attr, field, value, msg in (("NORMAL", "componentType", 5123, "NORMAL must be FLOAT"), ("POSITION", "type", "VEC2", "POSITION must be FLOAT"),
                                    ("TEXCOORD_0", "type", "VEC3", "TEXCOORD_0 must be FLOAT")):
        g = json.loads(json.dumps(good))
        if attr not in prim["attributes"]:
            continue
        g["accessors"][prim["attributes"][attr]][field] = value
        json.dump(g, open(path, "w"))
        r = subprocess.run([CLI, "render_option.json", "--dump-scene", "x.bin", "--quiet"], cwd=str(tmp_path), capture_output=True, text=True)
        assert r.returncode == 1 and msg in r.stderr, (attr, r.stderr)
    # a mesh whose only primitive has no triangles: the node contributes no instance, the rest of the scene loads
    g = json.loads(json.dumps(good))
    n_inst = sum(1 for n in g["nodes"] if "mesh" in n)
    empty_acc = json.loads(json.dumps(g["accessors"][prim["indices"]])); empty_acc["count"] = 0
    g["accessors"].append(empty_acc)
    g["meshes"].append({"primitives": [{"attributes": dict(prim["attributes"]), "indices": len(g["accessors"]) - 1, "material": 0}]})
    g["nodes"].insert(1, {"mesh": len(g["meshes"]) - 1, "name": "empty"})
    if "scenes" in g:
        g["scenes"][0]["nodes"] = list(range(len(g["nodes"])))
    for a in g.get("animations", []):
        for ch in a["channels"]:
            if ch["target"]["node"] >= 1:
                ch["target"]["node"] += 1
    json.dump(g, open(path, "w"))
    run_cli(["render_option.json", "--dump-scene", "scene.bin", "--quiet"], str(tmp_path))
    dump = read_dump(str(tmp_path / "scene.bin"))
    assert dump["prim_offsets"].size == n_inst and np.all(np.diff(dump["prim_offsets"].astype(np.int64)) > 0)
    assert dump["indices"].size == scene.indices.size
