#!/usr/bin/env python
"""Writes a SceneData (henjou-renderer_b200/scene.py) as a glTF 2.0 scene (.gltf + .bin + PNG textures) that the
loaders — the C++ host (henjou-renderer_b200/host/hj_gltf.h) and tools/gltf_to_npz.py — read back into the same arrays.
Used by the tests to feed the headless CLI on the GPU box, where the reference's own glTF files do not exist.

    python tools/npz_to_gltf.py scene.npz out_dir name        -> out_dir/name.gltf, out_dir/name.bin, out_dir/name_texN.png
"""
import json
import os
import struct
import sys
import zlib

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def write_png(path, rgba):
    """8-bit RGBA, filter type 0 (zlib only: no imaging library needed)."""
    h, w = rgba.shape[:2]
    raw = b"".join(b"\x00" + np.ascontiguousarray(rgba[y], np.uint8).tobytes() for y in range(h))

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xffffffff)
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 6, 0, 0, 0)) +
                chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def read_png(path):
    """Decoder for the PNGs the CLI writes (RGBA8, any filter)."""
    d = open(path, "rb").read()
    assert d[:8] == b"\x89PNG\r\n\x1a\n"
    p, idat, w, h = 8, b"", 0, 0
    while p < len(d):
        n, tag = struct.unpack(">I4s", d[p:p + 8])
        body = d[p + 8:p + 8 + n]
        if tag == b"IHDR":
            w, h, depth, ctype = struct.unpack(">IIBB", body[:10])
            assert depth == 8 and ctype == 6
        elif tag == b"IDAT":
            idat += body
        p += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, w * 4 + 1)
    out = np.zeros((h, w * 4), np.int32)
    for y in range(h):
        ft, row = raw[y, 0], raw[y, 1:].astype(np.int32)
        up = out[y - 1] if y else np.zeros(w * 4, np.int32)
        if ft == 0:
            out[y] = row
        elif ft == 2:
            out[y] = (row + up) & 255
        else:
            cur = out[y]
            for x in range(w * 4):
                a = cur[x - 4] if x >= 4 else 0
                b, c = up[x], (up[x - 4] if x >= 4 else 0)
                if ft == 1:
                    pr = a
                elif ft == 3:
                    pr = (a + b) >> 1
                else:
                    pa, pb, pc = abs(b - c), abs(a - c), abs(a + b - 2 * c)
                    pr = a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)
                cur[x] = (row[x] + pr) & 255
    return out.astype(np.uint8).reshape(h, w, 4)


def export_gltf(scene, out_dir, name, fov_fallback=None):
    os.makedirs(out_dir, exist_ok=True)
    blob = bytearray()
    views, accessors = [], []

    def add(arr, ctype, typ, with_minmax=False):
        arr = np.ascontiguousarray(arr)
        while len(blob) % 4:
            blob.append(0)
        views.append(dict(buffer=0, byteOffset=len(blob), byteLength=arr.nbytes))
        blob.extend(arr.tobytes())
        acc = dict(bufferView=len(views) - 1, componentType=ctype, count=int(arr.shape[0]), type=typ)
        if with_minmax:
            acc["min"] = [float(x) for x in np.atleast_1d(arr.min(0))]
            acc["max"] = [float(x) for x in np.atleast_1d(arr.max(0))]
        accessors.append(acc)
        return len(accessors) - 1

    g = dict(asset=dict(version="2.0", generator="henjou-renderer_b200 tools/npz_to_gltf.py"), scene=0)
    images, textures = [], []
    for i, (img, _srgb) in enumerate(scene.textures):
        uri = "%s_tex%d.png" % (name, i)
        write_png(os.path.join(out_dir, uri), np.asarray(img, np.uint8))
        images.append(dict(uri=uri)); textures.append(dict(source=i))
    mats = []
    for m in scene.materials:
        pbr = dict(baseColorFactor=[float(np.float32(x)) for x in m["basecolor"]] + [1.0], roughnessFactor=float(np.float32(m["roughness"])),
                   metallicFactor=float(np.float32(m["metallic"])))
        if m["basecolor_tex"] >= 0:
            pbr["baseColorTexture"] = dict(index=int(m["basecolor_tex"]))
        if m["roughness_tex"] >= 0:
            pbr["metallicRoughnessTexture"] = dict(index=int(m["roughness_tex"]))
        jm = dict(pbrMetallicRoughness=pbr)
        ext = {}
        em = np.asarray(m["emission"], np.float32)
        if float(em.max()) > 0:
            # emissiveFactor <= 1 per the spec; the rest goes to KHR_materials_emissive_strength (power of two: exact in float)
            s = 1.0
            while float(em.max()) / s > 1.0:
                s *= 2.0
            jm["emissiveFactor"] = [float(np.float32(x / np.float32(s))) for x in em]
            if s != 1.0:
                ext["KHR_materials_emissive_strength"] = dict(emissiveStrength=s)
        if m.get("normal_tex", -1) >= 0:
            jm["normalTexture"] = dict(index=int(m["normal_tex"]))
        if m["clearcoat"] != 0:
            ext["KHR_materials_clearcoat"] = dict(clearcoatFactor=float(np.float32(m["clearcoat"])))
        if m["sheen"] != 0:
            ext["KHR_materials_sheen"] = dict(sheenRoughnessFactor=float(np.float32(m["sheen"])))
        if m["transmission"] != 0:
            ext["KHR_materials_transmission"] = dict(transmissionFactor=float(np.float32(m["transmission"])))
        if m["ior"] != 1.0:
            ext["KHR_materials_ior"] = dict(ior=float(np.float32(m["ior"])))
        if m["is_thinfilm"]:
            ext["ThinFilm"] = dict(is_ThinFilm=True)
        if ext:
            jm["extensions"] = ext
        mats.append(jm)
    nodes, meshes, channels, samplers = [], [], [], []

    def node_trs(node, anim):
        node["translation"] = [float(x) for x in anim.t_val[0]]
        node["rotation"] = [float(x) for x in anim.r_val[0]]
        node["scale"] = [float(x) for x in anim.s_val[0]]

    def add_channels(ni, anim):
        for path, keys, vals, typ in (("translation", anim.t_key, anim.t_val, "VEC3"), ("rotation", anim.r_key, anim.r_val, "VEC4"),
                                      ("scale", anim.s_key, anim.s_val, "VEC3")):
            if len(keys) > 1:
                ka = add(np.asarray(keys[1:], np.float32), 5126, "SCALAR", True)
                va = add(np.asarray(vals[1:], np.float32), 5126, typ)
                samplers.append(dict(input=ka, output=va, interpolation="LINEAR"))
                channels.append(dict(sampler=len(samplers) - 1, target=dict(node=ni, path=path)))

    ntri = scene.n_triangles
    bounds = list(scene.prim_offsets.astype(np.int64)) + [ntri]
    from henjou_renderer_b200.scene import NodeAnimation
    for i in range(scene.n_instances):
        t0, t1 = int(bounds[i]), int(bounds[i + 1])
        prims = []
        mids = scene.material_ids[t0:t1]
        cuts = [0] + [int(k) + 1 for k in np.nonzero(mids[1:] != mids[:-1])[0]] + [t1 - t0]
        for a, b in zip(cuts[:-1], cuts[1:]):
            if b <= a:
                continue
            v0, v1 = 3 * (t0 + a), 3 * (t0 + b)
            pos = add(scene.vertices[v0:v1].astype(np.float32), 5126, "VEC3", True)
            nrm = add(scene.normals[v0:v1].astype(np.float32), 5126, "VEC3")
            uv = add(scene.texcoords[v0:v1].astype(np.float32), 5126, "VEC2")
            idx = add(np.arange(v1 - v0, dtype=np.uint32), 5125, "SCALAR")
            prims.append(dict(attributes=dict(POSITION=pos, NORMAL=nrm, TEXCOORD_0=uv), indices=idx, material=int(mids[a])))
        meshes.append(dict(primitives=prims))
        node = dict(mesh=i, name="inst%d" % i)
        anim = scene.animations[i] if i < len(scene.animations) and scene.animations[i] is not None else NodeAnimation()
        node_trs(node, anim)
        nodes.append(node)
        add_channels(len(nodes) - 1, anim)
    if scene.camera is not None and scene.camera.get("animation") is not None:
        cam = scene.camera
        g["cameras"] = [dict(type="perspective", perspective=dict(yfov=float(np.float32(cam["fov"])), znear=0.01, aspectRatio=1.0))]
        node = dict(camera=0, name="camera")
        node_trs(node, cam["animation"])
        nodes.append(node)
        add_channels(len(nodes) - 1, cam["animation"])
    g["nodes"] = nodes
    g["scenes"] = [dict(nodes=list(range(len(nodes))))]
    g["meshes"] = meshes
    g["materials"] = mats
    if images:
        g["images"], g["textures"] = images, textures
    if channels:
        g["animations"] = [dict(channels=channels, samplers=samplers)]
    g["accessors"], g["bufferViews"] = accessors, views
    g["buffers"] = [dict(uri=name + ".bin", byteLength=len(blob))]
    with open(os.path.join(out_dir, name + ".bin"), "wb") as f:
        f.write(bytes(blob))
    with open(os.path.join(out_dir, name + ".gltf"), "w") as f:
        json.dump(g, f)
    return os.path.join(out_dir, name + ".gltf")


def write_render_option(path, gltf_dir, gltf_name, image_name, width, height, spp, sky=(0.8, 0.8, 0.8), fps=24, start_frame=1,
                        end_frame=2, camera_position=(0, 1, -7), camera_direction=(0, 0, 1), camera_fov=45.0, lut_path="./LUT/Thin_Film_LUT.png",
                        allow_camera_animation=True, integrator=None, render_mode="Default", time_limit=5.0, ibl_path=None, ibl_intensity=1.0):
    """A render_option.json with the reference's schema (HenjouRenderer/render_option.json)."""
    j = dict(Image=dict(image_width=width, image_height=height, image_name=image_name, image_directory="./", max_spp=spp),
             Render_mode=render_mode, GLTF_file=dict(gltf_filepath=gltf_dir, gltf_filename=gltf_name),
             Camera=dict(allow_camera_animation=allow_camera_animation, camera_position=list(camera_position),
                         camera_direction=list(camera_direction), camera_fov=camera_fov),
             PTX_File=dict(ptxfile_path="../lib/ptx/Release/HenjouRenderer_generated_henjouRendererCU.cu.optixir"),
             Animation=dict(fps=fps, start_frame=start_frame, end_frame=end_frame, time_limit=time_limit),
             Sky=dict(IBL_path=ibl_path or "./HDRI/none.hdr", IBL_intensity=ibl_intensity, use_IBL=ibl_path is not None, scene_sky_default=list(sky)),
             Option=dict(use_date=True, save_renderOption=False), LUT=dict(LUT_path=lut_path))
    if integrator:
        j["Integrator"] = integrator
    with open(path, "w") as f:
        json.dump(j, f, indent=1)


if __name__ == "__main__":
    import henjou_renderer_b200 as hj
    sc = hj.load_npz(sys.argv[1])
    print("wrote", export_gltf(sc, sys.argv[2], sys.argv[3]))
