"""`.npz` (de)serialisation of SceneData — the fixture format under tests/golden/ (scenes converted from the
reference's glTF files by tools/gltf_to_npz.py, which needs /root/reference and therefore runs only in the
build container)."""
import json
import numpy as np

from .scene import SceneData, NodeAnimation


def _anim_to_json(a):
    if a is None:
        return None
    return dict(t_key=[float(k) for k in a.t_key], t_val=[[float(x) for x in v] for v in a.t_val],
                r_key=[float(k) for k in a.r_key], r_val=[[float(x) for x in v] for v in a.r_val],
                s_key=[float(k) for k in a.s_key], s_val=[[float(x) for x in v] for v in a.s_val])


def _anim_from_json(j):
    if j is None:
        return None
    a = NodeAnimation()
    a.t_key, a.r_key, a.s_key = j["t_key"], j["r_key"], j["s_key"]
    a.t_val = [np.asarray(v, np.float32) for v in j["t_val"]]
    a.r_val = [np.asarray(v, np.float32) for v in j["r_val"]]
    a.s_val = [np.asarray(v, np.float32) for v in j["s_val"]]
    return a


def save_npz(scene, path):
    meta = dict(materials=[{k: (list(map(float, v)) if isinstance(v, (tuple, list, np.ndarray)) else v) for k, v in m.items()}
                           for m in scene.materials],
                tex_srgb=[int(s) for _, s in scene.textures],
                animations=[_anim_to_json(a) for a in scene.animations],
                camera=None if scene.camera is None else dict(
                    position=[float(x) for x in scene.camera["position"]], direction=[float(x) for x in scene.camera["direction"]],
                    fov=float(scene.camera["fov"]), animation=_anim_to_json(scene.camera.get("animation"))))
    arrays = dict(vertices=scene.vertices, normals=scene.normals, texcoords=scene.texcoords, indices=scene.indices,
                  material_ids=scene.material_ids, prim_offsets=scene.prim_offsets, light_prim_ids=scene.light_prim_ids,
                  light_prim_emission=scene.light_prim_emission, ibl=scene.ibl, meta=np.frombuffer(json.dumps(meta).encode(), np.uint8))
    for i, (img, _) in enumerate(scene.textures):
        arrays["tex%d" % i] = img
    if scene.lut is not None:
        arrays["lut"] = scene.lut
    np.savez_compressed(path, **arrays)


def load_npz(path):
    z = np.load(path)
    s = SceneData()
    for k in ("vertices", "normals", "texcoords", "indices", "material_ids", "prim_offsets", "light_prim_ids",
              "light_prim_emission", "ibl"):
        setattr(s, k, z[k])
    meta = json.loads(bytes(z["meta"]).decode())
    s.materials = [{k: (tuple(v) if isinstance(v, list) else v) for k, v in m.items()} for m in meta["materials"]]
    s.textures = [(z["tex%d" % i], srgb) for i, srgb in enumerate(meta["tex_srgb"])]
    s.lut = z["lut"] if "lut" in z.files else None
    s.animations = [_anim_from_json(a) for a in meta["animations"]]
    cam = meta["camera"]
    if cam is not None:
        s.camera = dict(position=np.asarray(cam["position"], np.float32), direction=np.asarray(cam["direction"], np.float32),
                        fov=cam["fov"], animation=_anim_from_json(cam["animation"]))
    return s
