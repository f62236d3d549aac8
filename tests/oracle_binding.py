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
    return L


_libs = {}


def lib(kind="restatement"):
    """kind: 'restatement' (oracle/libhjo.so) or 'reference' (oracle/_ref/libhjref.so; None if not built)."""
    if kind not in _libs:
        path = os.path.join(ORACLE_DIR, "libhjo.so" if kind == "restatement" else "_ref/libhjref.so")
        if not os.path.exists(path):
            try:
                build_oracle()
            except Exception:
                pass
        _libs[kind] = _load(path) if os.path.exists(path) else None
    return _libs[kind]


def fp(a):
    return a.ctypes.data_as(abi.f32p)


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

    def close(self):
        if self.ctx:
            self.L.hjo_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
