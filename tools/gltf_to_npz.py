#!/usr/bin/env python
"""Fixture generator: converts a glTF scene of the reference into the `.npz` SceneData fixture format.

Follows the reference loader's semantics (/root/reference/include/loader/gltfloader.h:1068-1601): one instance
per mesh node, de-indexed triangles, identity index buffer, material extensions (KHR clearcoat / sheen /
transmission / ior / emissive_strength, custom ThinFilm), key-0 injection of the static TRS, and the
sampler-indexed-by-channel quirk (:1541-1542). Runs only where /root/reference exists; its outputs are
committed under tests/golden/ so the GPU box never needs the reference.

usage: python tools/gltf_to_npz.py <scene.gltf> <out.npz> [--sky r g b]
"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from henjou_renderer_b200.scene import SceneData, NodeAnimation, make_material   # noqa: E402
from henjou_renderer_b200.sceneio import save_npz                                  # noqa: E402

_COMP = {5120: np.int8, 5121: np.uint8, 5122: np.int16, 5123: np.uint16, 5124: np.int32, 5125: np.uint32, 5126: np.float32}
_NCOMP = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4}


def _accessor(g, buffers, idx):
    a = g["accessors"][idx]
    bv = g["bufferViews"][a["bufferView"]]
    dt = np.dtype(_COMP[a["componentType"]])
    nc = _NCOMP[a["type"]]
    off = bv.get("byteOffset", 0) + a.get("byteOffset", 0)
    stride = bv.get("byteStride", 0) or dt.itemsize * nc
    buf = buffers[bv["buffer"]]
    out = np.zeros((a["count"], nc), dt)
    for i in range(a["count"]):
        out[i] = np.frombuffer(buf, dt, nc, off + i * stride)
    return out


def load_gltf(path, allow_camera_animation=True):
    from PIL import Image
    base = os.path.dirname(path)
    g = json.load(open(path))
    buffers = [open(os.path.join(base, b["uri"]), "rb").read() for b in g["buffers"]]
    s = SceneData()
    known = {}

    def tex(index, srgb):
        uri = g["images"][g["textures"][index]["source"]]["uri"]
        if uri not in known:                                           # loader/texture_load.h:7 cache by name
            img = np.asarray(Image.open(os.path.join(base, uri)).convert("RGBA"), np.uint8)
            known[uri] = len(s.textures)
            s.textures.append((img, int(srgb)))
        return known[uri]

    for m in g.get("materials", []):
        pbr = m.get("pbrMetallicRoughness", {})
        kw = dict(basecolor=tuple(np.float32(pbr.get("baseColorFactor", [1, 1, 1, 1])[:3]).tolist()),
                  roughness=float(np.float32(pbr.get("roughnessFactor", 1.0))),
                  metallic=float(np.float32(pbr.get("metallicFactor", 1.0))))
        kw["basecolor_tex"] = tex(pbr["baseColorTexture"]["index"], 1) if "baseColorTexture" in pbr else -1
        kw["roughness_tex"] = tex(pbr["metallicRoughnessTexture"]["index"], 0) if "metallicRoughnessTexture" in pbr else -1
        kw["metallic_tex"] = kw["roughness_tex"]                       # gltfloader.h:1156
        em = np.float32(m.get("emissiveFactor", [0, 0, 0]))
        is_light_before_strength = float(em[0]) + float(em[1]) + float(em[2]) > 0.0
        kw["normal_tex"] = tex(m["normalTexture"]["index"], 0) if "normalTexture" in m else -1
        ext = m.get("extensions", {})
        if "clearcoatFactor" in ext.get("KHR_materials_clearcoat", {}):
            kw["clearcoat"] = float(np.float32(ext["KHR_materials_clearcoat"]["clearcoatFactor"]))
        if "sheenRoughnessFactor" in ext.get("KHR_materials_sheen", {}):   # sic, gltfloader.h:1211
            kw["sheen"] = float(np.float32(ext["KHR_materials_sheen"]["sheenRoughnessFactor"]))
        if "transmissionFactor" in ext.get("KHR_materials_transmission", {}):
            kw["transmission"] = float(np.float32(ext["KHR_materials_transmission"]["transmissionFactor"]))
        if "ior" in ext.get("KHR_materials_ior", {}):
            kw["ior"] = float(np.float32(ext["KHR_materials_ior"]["ior"]))
        if "emissiveStrength" in ext.get("KHR_materials_emissive_strength", {}):
            em = (em.astype(np.float64) * ext["KHR_materials_emissive_strength"]["emissiveStrength"]).astype(np.float32)
        if "is_ThinFilm" in ext.get("ThinFilm", {}):                    # presence, not value (gltfloader.h:1250-1256)
            kw["is_thinfilm"] = 1
        kw["emission"] = tuple(em.tolist())
        mat = make_material(**kw)
        mat["is_light"] = int(is_light_before_strength)
        s.materials.append(mat)

    node_anims = []
    for n in g["nodes"]:
        node_anims.append(NodeAnimation(n.get("translation", (0, 0, 0)), n.get("rotation", (0, 0, 0, 1)), n.get("scale", (1, 1, 1))))
    V, N, T, MID, PO, inst_nodes = [], [], [], [], [], []
    ntri = 0
    for ni, n in enumerate(g["nodes"]):
        if "mesh" in n:
            PO.append(ntri)
            inst_nodes.append(ni)
            for prim in g["meshes"][n["mesh"]]["primitives"]:
                idx = _accessor(g, buffers, prim["indices"]).reshape(-1).astype(np.int64)
                at = prim["attributes"]
                pos = _accessor(g, buffers, at["POSITION"]).astype(np.float32)
                nt = idx.shape[0] // 3
                idx = idx[: nt * 3]
                v = pos[idx]
                if "NORMAL" in at:
                    nn = _accessor(g, buffers, at["NORMAL"]).astype(np.float32)[idx]
                else:
                    tri = v.reshape(nt, 3, 3)
                    gn = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0]).astype(np.float32)
                    gn = gn * (np.float32(1.0) / np.sqrt(np.sum(gn * gn, axis=1, dtype=np.float32)))[:, None]
                    nn = np.repeat(gn, 3, axis=0)
                tc = _accessor(g, buffers, at["TEXCOORD_0"]).astype(np.float32)[idx] if "TEXCOORD_0" in at else np.zeros((nt * 3, 2), np.float32)
                V.append(v); N.append(nn); T.append(tc)
                MID.append(np.full(nt, prim["material"], np.uint32))
                ntri += nt
        elif "camera" in n and allow_camera_animation:
            s.camera = dict(position=np.zeros(3, np.float32), direction=np.array([0, 0, -1], np.float32),
                            fov=float(np.float32(g["cameras"][n["camera"]]["perspective"]["yfov"])), animation=node_anims[ni])
    s.vertices = np.concatenate(V).astype(np.float32)
    s.normals = np.concatenate(N).astype(np.float32)
    s.texcoords = np.concatenate(T).astype(np.float32)
    s.material_ids = np.concatenate(MID)
    s.indices = np.arange(3 * ntri, dtype=np.uint32)
    s.prim_offsets = np.asarray(PO, np.uint32)
    s.finalize_lights()
    for anim in g.get("animations", []):
        for i, ch in enumerate(anim["channels"]):
            smp = anim["samplers"][i]                                   # sic: channel index, gltfloader.h:1541-1542
            keys = _accessor(g, buffers, smp["input"]).reshape(-1)
            data = _accessor(g, buffers, smp["output"]).astype(np.float32)
            a = node_anims[ch["target"]["node"]]
            path = ch["target"]["path"]
            for k in range(data.shape[0]):
                if path == "translation":
                    a.t_val.append(data[k]); a.t_key.append(float(keys[k]))
                elif path == "rotation":
                    a.r_val.append(data[k]); a.r_key.append(float(keys[k]))
                elif path == "scale":
                    a.s_val.append(data[k]); a.s_key.append(float(keys[k]))
    s.animations = [node_anims[ni] for ni in inst_nodes]
    return s


if __name__ == "__main__":
    src, dst = sys.argv[1], sys.argv[2]
    sc = load_gltf(src)
    if "--sky" in sys.argv:
        k = sys.argv.index("--sky")
        sc.set_sky([float(x) for x in sys.argv[k + 1:k + 4]])
    save_npz(sc, dst)
    print("wrote", dst, "tris", sc.n_triangles, "instances", sc.n_instances, "materials", len(sc.materials),
          "lights", sc.light_prim_ids.shape[0], "textures", len(sc.textures))
