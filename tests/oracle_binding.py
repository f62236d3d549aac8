"""ctypes binding of the CPU oracle (oracle/libhjo.so, oracle/_ref/libhjref.so). TEST INFRASTRUCTURE:
imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
import ctypes as C
import os
import subprocess

import numpy as np

import henjou_renderer_b200 as hj
from henjou_renderer_b200 import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")


class hjo_render_out(C.Structure):
    _fields_ = [("color", abi.f32p), ("albedo", abi.f32p), ("normal", abi.f32p), ("rgba8", abi.u8p), ("accum", abi.f32p),
                ("path_radiance", abi.f32p), ("path_ids", abi.u32p),
                ("rays_extend", C.c_uint64), ("rays_shadow", C.c_uint64), ("traced_extend", C.c_uint64), ("traced_shadow", C.c_uint64),
                ("seconds", C.c_double)]


class hjo_kat_surface(C.Structure):
    _fields_ = [("basecolor", C.c_float * 3), ("metallic", C.c_float), ("roughness", C.c_float), ("sheen", C.c_float),
                ("clearcoat", C.c_float), ("ior", C.c_float), ("is_specular", C.c_int32), ("is_thinfilm", C.c_int32)]


def build_oracle():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True)


def _load(path):
    L = C.CDLL(path)
    L.hjo_create.restype = C.c_void_p
    L.hjo_destroy.argtypes = [C.c_void_p]
    L.hjo_scene_upload.argtypes = [C.c_void_p, C.POINTER(abi.hj_scene_desc)]
    L.hjo_accel_build.argtypes = [C.c_void_p, abi.f32p, abi.f32p, C.c_uint32, C.c_int]
    L.hjo_render.argtypes = [C.c_void_p, C.POINTER(abi.hj_frame_params), C.c_int, C.c_uint32, C.c_uint32, C.POINTER(hjo_render_out)]
    L.hjo_trace.argtypes = [C.c_void_p, abi.f32p, C.c_uint64, C.c_int, abi.u32p, abi.f32p]
    for n in ("hjo_kat_xxhash32", "hjo_kat_xxhash32_u4", "hjo_kat_permute", "hjo_kat_cmj2d"):
        getattr(L, n).restype = C.c_uint32
    L.hjo_kat_cmj2d.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, abi.f32p]
    L.hjo_kat_randfloat.restype = C.c_float
    L.hjo_kat_light_pdf.restype = C.c_float
    L.hjo_kat_light_pdf.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
    L.hjo_kat_light_sample.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, abi.f32p]
    L.hjo_kat_bsdf_sample.argtypes = [C.POINTER(hjo_kat_surface), abi.f32p, C.c_uint64, C.c_uint32, C.c_uint32, abi.f32p]
    L.hjo_kat_bsdf_eval.argtypes = [C.POINTER(hjo_kat_surface), abi.f32p, abi.f32p, abi.f32p]
    L.hjo_kat_ideal_glass.argtypes = [C.c_float, abi.f32p, C.c_uint64, C.c_uint32, C.c_uint32, abi.f32p]
    L.hjo_kat_set_lut_const.argtypes = [C.c_int, C.c_float, C.c_float, C.c_float]
    L.hjo_kat_cosine.argtypes = [C.c_float, C.c_float, abi.f32p, abi.f32p]
    L.hjo_kat_refract.argtypes = [abi.f32p, abi.f32p, C.c_float, C.c_float, abi.f32p]
    L.hjo_kat_onb.argtypes = [abi.f32p, abi.f32p, abi.f32p]
    L.hjo_kat_tonemap.argtypes = [C.c_int, abi.f32p, abi.f32p]
    L.hjo_kat_to_srgb8.argtypes = [abi.f32p, abi.u8p]
    L.hjo_path_events.argtypes = [C.c_void_p, C.POINTER(abi.hj_frame_params), C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint32, abi.f32p]
    L.hjo_kat_libm.argtypes = [C.c_int, abi.f32p, abi.f32p, abi.f32p, C.c_uint64]
    return L


_libs = {}


PATHS = {"restatement": "libhjo.so", "glibc": "libhjo_glibc.so", "reference": "_ref/libhjref.so"}
LIBM_FN = {"sinf": 0, "cosf": 1, "acosf": 2, "logf": 3, "expf": 4, "powf": 5, "atan2f": 6}
EVENT_DTYPE = np.dtype([("kind", np.uint32), ("outcome", np.uint32), ("a", np.float32), ("b", np.float32)])
EVENT_NAMES = {1: "russian roulette", 2: "Disney lobe: diffuse?", 3: "Disney lobe: specular?", 4: "Disney wi below horizon", 5: "glass: inside?",
               6: "glass: Fresnel reflect?", 7: "glass: total internal reflection", 8: "EC-GGX: wr.y > 0.9999", 9: "EC-GGX: wr.y < -0.9999",
               10: "EC-GGX: grazing", 11: "EC-GGX: escape (U > 1 - G1)", 12: "EC-GGX: wi below horizon", 13: "light pick", 20: "closest hit",
               21: "shadow ray"}


def lib(kind="restatement"):
    """kind: 'restatement' (oracle/libhjo.so: the restatement with the product's shared hj_libm.h, bit-comparable to the
    kernels), 'glibc' (oracle/libhjo_glibc.so: the same restatement on glibc's libm, the bridge to the reference build) or
    'reference' (oracle/_ref/libhjref.so: the reference's own headers; None if not built)."""
    if kind not in _libs:
        path = os.path.join(ORACLE_DIR, PATHS[kind])
        if not os.path.exists(path):
            try:
                build_oracle()
            except Exception:
                pass
        _libs[kind] = _load(path) if os.path.exists(path) else None
    return _libs[kind]


def fp(a):
    return a.ctypes.data_as(abi.f32p)


def libm(kind, fn, a, b=None):
    """Evaluates one elementary function (LIBM_FN) as the oracle build `kind` does, over float32 arrays."""
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b if b is not None else np.zeros_like(a), np.float32)
    out = np.empty_like(a)
    assert lib(kind).hjo_kat_libm(LIBM_FN[fn], fp(a), fp(b), fp(out), a.size) == 0
    return out


class Oracle:
    def __init__(self, scene, kind="restatement", use_bvh=True, xforms=None, inv_xforms=None):
        self.L = lib(kind)
        if self.L is None:
            raise RuntimeError("oracle build '%s' not available" % kind)
        self.ctx = self.L.hjo_create()
        self.scene = scene
        desc = scene.to_desc()
        assert self.L.hjo_scene_upload(self.ctx, C.byref(desc)) == 0
        self.set_transforms(xforms, inv_xforms, use_bvh)

    def set_transforms(self, xforms=None, inv_xforms=None, use_bvh=True):
        xf = None if xforms is None else np.ascontiguousarray(xforms, np.float32)
        ixf = None if inv_xforms is None else np.ascontiguousarray(inv_xforms, np.float32)
        rc = self.L.hjo_accel_build(self.ctx, fp(xf) if xf is not None else None, fp(ixf) if ixf is not None else None,
                                    self.scene.n_instances, int(use_bvh))
        assert rc == 0, rc

    def render(self, fparams, threads=0, rows=(0, 0), debug=False):
        W, H = fparams.width, fparams.height
        ns = fparams.sample_end - fparams.sample_begin
        out = hjo_render_out()
        res = dict(color=np.zeros((H, W, 4), np.float32), albedo=np.zeros((H, W, 4), np.float32), normal=np.zeros((H, W, 4), np.float32),
                   rgba8=np.zeros((H, W, 4), np.uint8), accum=np.zeros((H, W, 4), np.float32))
        if debug:
            res["path_radiance"] = np.zeros((ns, H, W, 4), np.float32)
            res["path_ids"] = np.zeros((ns, H, W, 2), np.uint32)
            out.path_radiance = fp(res["path_radiance"])
            out.path_ids = res["path_ids"].ctypes.data_as(abi.u32p)
        out.color, out.albedo, out.normal, out.accum = fp(res["color"]), fp(res["albedo"]), fp(res["normal"]), fp(res["accum"])
        out.rgba8 = res["rgba8"].ctypes.data_as(abi.u8p)
        rc = self.L.hjo_render(self.ctx, C.byref(fparams), threads, rows[0], rows[1], C.byref(out))
        assert rc == 0, rc
        res.update(rays_extend=out.rays_extend, rays_shadow=out.rays_shadow, traced_extend=out.traced_extend,
                   traced_shadow=out.traced_shadow, seconds=out.seconds)
        return res

    def trace(self, rays, any_hit=False):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        n = rays.shape[0]
        prim = np.zeros(n, np.uint32)
        tuv = np.zeros((n, 3), np.float32)
        assert self.L.hjo_trace(self.ctx, fp(rays), n, int(any_hit), prim.ctypes.data_as(abi.u32p), fp(tuv)) == 0
        return prim, tuv

    def path_events(self, fparams, pixel, sample, cap=4096):
        """Decision trace of one path: (structured array of events, radiance[3])."""
        ev = np.zeros(cap, EVENT_DTYPE)
        rad = np.zeros(3, np.float32)
        n = self.L.hjo_path_events(self.ctx, C.byref(fparams), int(pixel), int(sample), ev.ctypes.data_as(C.c_void_p), cap, fp(rad))
        assert n >= 0, n
        return ev[:min(n, cap)], rad

    def close(self):
        if self.ctx:
            self.L.hjo_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
