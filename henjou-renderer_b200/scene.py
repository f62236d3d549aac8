"""Host-side scene container: the nine SceneData arrays of the reference
(/root/reference/include/renderer/scene.h:19-37) as numpy arrays, plus materials, textures, sky, LUT,
per-node TRS animation (include/renderer/animation.h:34-131) and the camera maths of the frame loop
(include/renderer/renderer.h:1145-1169). `to_desc()` lends the arrays to the C ABI without copying."""
import ctypes as C
import math
import numpy as np

from . import abi

MATERIAL_DEFAULTS = dict(
    basecolor=(1.0, 1.0, 1.0), basecolor_tex=-1, emission=(0.0, 0.0, 0.0), emission_tex=-1, is_light=0,
    metallic=1.0, metallic_tex=-1, roughness=1.0, roughness_tex=-1, specular=(0.0, 0.0, 0.0), sheen=0.0,
    clearcoat=0.0, clearcoat_tex=-1, bump_tex=-1, normal_tex=-1, transmission=0.0, ior=1.0,
    ideal_specular=0, is_thinfilm=0)


def make_material(**kw):
    """A material dict with the loader's defaults (gltfloader.h:1125-1263 + tinygltf factor defaults).
    is_light / ideal_specular are derived exactly as the loader derives them."""
    m = dict(MATERIAL_DEFAULTS)
    m.update(kw)
    e = np.asarray(m["emission"], np.float32)
    m["is_light"] = int(float(e[0]) + float(e[1]) + float(e[2]) > 0.0)                 # gltfloader.h:1162-1167
    m["ideal_specular"] = int(m["roughness"] == 0 and m["transmission"] > 0)            # gltfloader.h:1260-1263
    return m


class NodeAnimation:
    """TRS key tracks of one node; key 0 is the node's static TRS (gltfloader.h:1313-1343)."""

    def __init__(self, t=(0, 0, 0), r=(0, 0, 0, 1), s=(1, 1, 1)):
        self.t_key, self.t_val = [0.0], [np.asarray(t, np.float32)]
        self.r_key, self.r_val = [0.0], [np.asarray(r, np.float32)]
        self.s_key, self.s_val = [0.0], [np.asarray(s, np.float32)]

    @staticmethod
    def _interp(vals, keys, time):
        # Animation::animationInterpolate, animation.h:39-67 (always linear, :69-79)
        if len(keys) == 1 or time < 0:
            return vals[0]
        first, ln = 0, len(keys)
        while ln > 0:
            half = ln >> 1
            mid = first + half
            if keys[mid] <= time:
                first = mid + 1
                ln -= half + 1
            else:
                ln = half
        off = first - 1
        if len(keys) - 1 <= off:
            return vals[len(keys) - 1]
        f32 = np.float32
        delta = f32(f32(time) - f32(keys[off])) / f32(f32(keys[off + 1]) - f32(keys[off]))
        return (vals[off] * f32(f32(1.0) - delta) + vals[off + 1] * delta).astype(np.float32)

    def trs(self, time):
        return (self._interp(self.t_val, self.t_key, time), self._interp(self.r_val, self.r_key, time),
                self._interp(self.s_val, self.s_key, time))

    @staticmethod
    def rotate_affine(q):
        # common/matrix.h rotateAffine: float products of (2.0*q.a*q.b) computed in double then rounded
        x, y, z, w = [float(v) for v in q]
        f = np.float32
        xy, xz, xw = f(2.0 * x * y), f(2.0 * x * z), f(2.0 * x * w)
        yz, yw, zw, ww = f(2.0 * y * z), f(2.0 * y * w), f(2.0 * z * w), f(2.0 * w * w)
        qx, qy, qz = f(x), f(y), f(z)
        m = np.eye(4, dtype=np.float32)
        m[0, :3] = [ww + f(2.0) * qx * qx - f(1.0), xy - zw, xz + yw]
        m[1, :3] = [xy + zw, ww + f(2.0) * qy * qy - f(1.0), yz - xw]
        m[2, :3] = [xz - yw, yz + xw, ww + f(2.0) * qz * qz - f(1.0)]
        return m

    def affine(self, time):
        """Animation::getAnimationAffine, animation.h:81-94: T * R * S in float32."""
        t, r, s = self.trs(time)
        T = np.eye(4, dtype=np.float32); T[:3, 3] = t
        S = np.diag(np.array([s[0], s[1], s[2], 1.0], np.float32))
        return _matmul_f32(_matmul_f32(T, self.rotate_affine(r)), S)

    def rotate_only(self, time):
        _, r, _ = self.trs(time)
        return self.rotate_affine(r)


def _matmul_f32(a, b):
    # common/matrix.h operator*: left-to-right float32 accumulation
    out = np.zeros((4, 4), np.float32)
    for j in range(4):
        for i in range(4):
            acc = np.float32(a[j, 0] * b[0, i])
            for k in (1, 2, 3):
                acc = np.float32(acc + np.float32(a[j, k] * b[k, i]))
            out[j, i] = acc
    return out


class SceneData:
    def __init__(self):
        self.vertices = np.zeros((0, 3), np.float32)
        self.normals = np.zeros((0, 3), np.float32)
        self.texcoords = np.zeros((0, 2), np.float32)
        self.indices = np.zeros((0,), np.uint32)
        self.material_ids = np.zeros((0,), np.uint32)
        self.prim_offsets = np.zeros((0,), np.uint32)
        self.materials = []                 # list of dicts (make_material)
        self.textures = []                  # list of (uint8[h,w,4], srgb flag)
        self.light_prim_ids = np.zeros((0,), np.uint32)
        self.light_prim_emission = np.zeros((0, 3), np.float32)
        self.ibl = np.zeros((1, 1, 4), np.float32)      # float4 image; 1x1 = constant sky
        self.lut = None                     # uint8[h,w,4] or None
        self.animations = []                # NodeAnimation per instance (instance i <-> mesh node i)
        self.camera = None                  # dict(position, direction, fov, animation=NodeAnimation|None)
        self._keep = []

    # ------------------------------------------------------------------ derived data
    @property
    def n_triangles(self):
        return int(self.material_ids.shape[0])

    @property
    def n_instances(self):
        return int(self.prim_offsets.shape[0])

    def finalize_lights(self):
        """light_prim_ids / light_prim_emission as the loader emits them (gltfloader.h:1496-1502)."""
        is_light = np.array([m["is_light"] for m in self.materials], bool)
        ids = np.nonzero(is_light[self.material_ids])[0].astype(np.uint32)
        em = np.array([self.materials[int(self.material_ids[i])]["emission"] for i in ids], np.float32).reshape(-1, 3)
        self.light_prim_ids, self.light_prim_emission = ids, em

    def set_sky(self, rgb):
        self.ibl = np.array([[[rgb[0], rgb[1], rgb[2], 0.0]]], np.float32)   # texture.h:58-65

    def set_sky_image(self, rgb):
        """Equirect sky image float[H,W,3] (or [H,W,4]) -> the float4 image of HDRTexture, w = 0 (texture.h:67-101)."""
        rgb = np.asarray(rgb, np.float32)
        img = np.zeros(rgb.shape[:2] + (4,), np.float32)
        img[..., :3] = rgb[..., :3]
        self.ibl = img

    def transforms(self, time=None):
        """Per-instance row-major 3x4 matrices and the rows of their inverses (renderer.h:257-291).
        time=None -> identity (cpySceneDataToDevice, renderer.h:197-255)."""
        n = self.n_instances
        xf = np.zeros((n, 3, 4), np.float32)
        inv = np.zeros((n, 3, 4), np.float32)
        for i in range(n):
            if time is None or i >= len(self.animations) or self.animations[i] is None:
                m = np.eye(4, dtype=np.float32)
            else:
                m = self.animations[i].affine(time)
            xf[i] = m[:3]
            inv[i] = np.linalg.inv(m.astype(np.float64))[:3].astype(np.float32)
        return xf, inv

    # ------------------------------------------------------------------ C ABI view
    def to_desc(self):
        d = abi.hj_scene_desc()
        keep = []

        def arr(a, dt, ptr):
            a = np.ascontiguousarray(a, dt)
            keep.append(a)
            return a.ctypes.data_as(ptr) if a.size else C.cast(None, ptr)

        d.vertices = arr(self.vertices, np.float32, abi.f32p)
        d.normals = arr(self.normals, np.float32, abi.f32p)
        d.texcoords = arr(self.texcoords, np.float32, abi.f32p)
        d.indices = arr(self.indices, np.uint32, abi.u32p)
        d.material_ids = arr(self.material_ids, np.uint32, abi.u32p)
        d.prim_offsets = arr(self.prim_offsets, np.uint32, abi.u32p)
        d.n_vertices = self.vertices.shape[0]
        d.n_triangles = self.n_triangles
        d.n_instances = self.n_instances
        mats = (abi.hj_material * max(1, len(self.materials)))()
        for i, m in enumerate(self.materials):
            for k, v in m.items():
                if k in ("basecolor", "emission", "specular"):
                    setattr(mats[i], k, (C.c_float * 3)(*[float(x) for x in v]))
                elif k in dict(abi.hj_material._fields_):
                    setattr(mats[i], k, v)
        keep.append(mats)
        d.materials = mats
        d.n_materials = len(self.materials)
        texs = (abi.hj_texture * max(1, len(self.textures)))()
        for i, (img, srgb) in enumerate(self.textures):
            img = np.ascontiguousarray(img, np.uint8)
            keep.append(img)
            texs[i].rgba8 = img.ctypes.data_as(abi.u8p)
            texs[i].height, texs[i].width = img.shape[0], img.shape[1]
            texs[i].srgb = int(srgb)
        keep.append(texs)
        d.textures = texs
        d.n_textures = len(self.textures)
        d.light_prim_ids = arr(self.light_prim_ids, np.uint32, abi.u32p)
        d.light_prim_emission = arr(self.light_prim_emission, np.float32, abi.f32p)
        d.n_light_prims = int(self.light_prim_ids.shape[0])
        d.ibl_rgba32f = arr(self.ibl, np.float32, abi.f32p)
        d.ibl_height, d.ibl_width = self.ibl.shape[0], self.ibl.shape[1]
        if self.lut is not None:
            d.lut_rgba8 = arr(self.lut, np.uint8, abi.u8p)
            d.lut_height, d.lut_width = self.lut.shape[0], self.lut.shape[1]
        self._keep = keep
        return d
