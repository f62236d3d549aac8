"""Host-side driver over the C ABI (libhenjou_b200.so): the calls the reference's `Renderer` makes around its
device code (/root/reference/include/renderer/renderer.h: cpySceneDataToDevice :197, updateIASMatrix :257,
launch :1229-1242, AOV read-back :1279), with the same order of operations. ctypes only: no torch types cross
the boundary. There is no CPU path: a missing library or a missing B200 raises."""
import ctypes as C
import os

import numpy as np

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HJ_LIB_PATH") or os.path.join(_HERE, "libhenjou_b200.so")   # HJ_LIB_PATH: kernel-tuning variants
_lib = None


class HenjouError(RuntimeError):
    pass


def load_library():
    """Loads libhenjou_b200.so (built in-tree by __graft_entry__.build() / csrc/Makefile) and types its exports."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise HenjouError("libhenjou_b200.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                          "there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    L.hj_abi_version.restype = C.c_int
    L.hj_create.argtypes = [C.POINTER(vp), C.c_int]
    L.hj_destroy.argtypes = [vp]
    L.hj_destroy.restype = None
    L.hj_last_error.argtypes = [vp]
    L.hj_last_error.restype = C.c_char_p
    L.hj_set_option.argtypes = [vp, C.c_char_p, C.c_int64]
    L.hj_scene_upload.argtypes = [vp, C.POINTER(abi.hj_scene_desc)]
    L.hj_accel_build.argtypes = [vp, abi.f32p, abi.f32p, C.c_uint32, C.c_int]
    L.hj_accel_build_async.argtypes = [vp, abi.f32p, abi.f32p, C.c_uint32, C.c_int]
    L.hj_accel_wait.argtypes = [vp]
    L.hj_render.argtypes = [vp, C.POINTER(abi.hj_frame_params)]
    L.hj_render_async.argtypes = [vp, C.POINTER(abi.hj_frame_params)]
    L.hj_sync.argtypes = [vp]
    L.hj_readback.argtypes = [vp, C.c_int, vp, C.c_size_t]
    L.hj_device_ptr.argtypes = [vp, C.c_int]
    L.hj_device_ptr.restype = vp
    L.hj_finalize.argtypes = [vp]
    L.hj_get_stats.argtypes = [vp, C.POINTER(abi.hj_stats)]
    L.hj_trace.argtypes = [vp, abi.f32p, C.c_uint64, C.c_int, abi.u32p, abi.f32p]
    L.hj_trace_noalloc.argtypes = [vp, abi.f32p, C.c_uint64, C.c_int, abi.u32p, abi.f32p]
    L.hj_post.argtypes = [vp, C.c_int]
    L.hj_libm_eval.argtypes = [vp, C.c_int, abi.f32p, abi.f32p, abi.f32p, C.c_uint64]
    L.hj_host_pin.argtypes = [vp, vp, C.c_size_t]
    L.hj_host_unpin.argtypes = [vp, vp]
    L.hj_comm_unique_id.argtypes = [vp]
    L.hj_comm_init.argtypes = [vp, vp, C.c_int, C.c_int]
    L.hj_comm_reduce.argtypes = [vp]
    L.hj_comm_scene_upload.argtypes = [vp, C.POINTER(abi.hj_scene_desc), C.c_int]
    _lib = L
    return L


class Renderer:
    """One context = one GPU. Usage mirrors Renderer::initializeAndRender (renderer.h:1053):
    upload(scene) -> per frame: build_accel(transforms) -> render(frame_params) -> readback(aov)."""

    def __init__(self, device=0):
        self.L = load_library()
        self.ctx = C.c_void_p()
        rc = self.L.hj_create(C.byref(self.ctx), device)
        if rc != 0:
            raise HenjouError("hj_create failed (%d): %s" % (rc, self.L.hj_last_error(None).decode()))
        self.scene = None
        self.fp = None

    def _check(self, rc, what):
        if rc != 0:
            raise HenjouError("%s failed (%d): %s" % (what, rc, self.L.hj_last_error(self.ctx).decode()))

    def set_option(self, key, value):
        self._check(self.L.hj_set_option(self.ctx, key.encode(), int(value)), "hj_set_option")

    def upload(self, scene):
        self.scene = scene
        desc = scene.to_desc()
        self._check(self.L.hj_scene_upload(self.ctx, C.byref(desc)), "hj_scene_upload")

    def build_accel(self, xforms=None, inv_xforms=None, mode=abi.ACCEL_BUILD, sync=True):
        """sync=False: hj_accel_build_async — enqueued on the build stream into the second buffer set (it may run under a frame
        in flight); the next render / trace / accel_wait completes it."""
        xf = None if xforms is None else np.ascontiguousarray(xforms, np.float32)
        ixf = None if inv_xforms is None else np.ascontiguousarray(inv_xforms, np.float32)
        p = lambda a: a.ctypes.data_as(abi.f32p) if a is not None else None
        if sync:
            self._check(self.L.hj_accel_build(self.ctx, p(xf), p(ixf), self.scene.n_instances, mode), "hj_accel_build")
        else:
            self._check(self.L.hj_accel_build_async(self.ctx, p(xf), p(ixf), self.scene.n_instances, mode), "hj_accel_build_async")

    def accel_wait(self):
        self._check(self.L.hj_accel_wait(self.ctx), "hj_accel_wait")

    def render(self, fparams, sync=True):
        self.fp = fparams
        if sync:
            self._check(self.L.hj_render(self.ctx, C.byref(fparams)), "hj_render")
        else:
            self._check(self.L.hj_render_async(self.ctx, C.byref(fparams)), "hj_render_async")

    def sync(self):
        self._check(self.L.hj_sync(self.ctx), "hj_sync")

    def post(self, mode=abi.POST_PASSTHROUGH):
        """OptixDenoiserManager::denoise + float4ConvertColor (renderer.h:1256-1283) -> AOV_OUTPUT / AOV_OUTPUT_RGBA8."""
        self._check(self.L.hj_post(self.ctx, int(mode)), "hj_post")
        self.post_scale = 2 if mode == abi.POST_DENOISE_UPSCALE2X else 1

    def readback(self, aov, out=None):
        W, H = self.fp.width, self.fp.height
        if aov in (abi.AOV_OUTPUT, abi.AOV_OUTPUT_RGBA8):
            W, H = W * getattr(self, "post_scale", 1), H * getattr(self, "post_scale", 1)
        ns = self.fp.sample_end - self.fp.sample_begin
        shape, dt = {abi.AOV_RGBA8: ((H, W, 4), np.uint8), abi.AOV_OUTPUT_RGBA8: ((H, W, 4), np.uint8), abi.AOV_PATH_RADIANCE: ((ns, H, W, 4), np.float32),
                     abi.AOV_PATH_IDS: ((ns, H, W, 2), np.uint32)}.get(aov, ((H, W, 4), np.float32))
        if out is None:
            out = np.empty(shape, dt)
        self._check(self.L.hj_readback(self.ctx, aov, out.ctypes.data_as(C.c_void_p), out.nbytes), "hj_readback")
        return out

    def pin(self, *arrays):
        """Page-locks numpy buffers handed to upload()/readback() (hj_host_pin); keep them alive until unpin()."""
        for a in arrays:
            if a is not None and a.nbytes:
                self._check(self.L.hj_host_pin(self.ctx, a.ctypes.data_as(C.c_void_p), a.nbytes), "hj_host_pin")

    def unpin(self, *arrays):
        for a in arrays:
            if a is not None and a.nbytes:
                self._check(self.L.hj_host_unpin(self.ctx, a.ctypes.data_as(C.c_void_p)), "hj_host_unpin")

    def device_ptr(self, aov):
        return self.L.hj_device_ptr(self.ctx, aov)

    def finalize(self):
        self._check(self.L.hj_finalize(self.ctx), "hj_finalize")

    def stats(self):
        s = abi.hj_stats()
        self._check(self.L.hj_get_stats(self.ctx, C.byref(s)), "hj_get_stats")
        return {k: getattr(s, k) for k, _ in abi.hj_stats._fields_}

    def trace(self, rays, any_hit=False):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        n = rays.shape[0]
        prim = np.zeros(n, np.uint32)
        tuv = np.zeros((n, 3), np.float32)
        self._check(self.L.hj_trace(self.ctx, rays.ctypes.data_as(abi.f32p), n, int(any_hit), prim.ctypes.data_as(abi.u32p),
                                    tuv.ctypes.data_as(abi.f32p)), "hj_trace")
        return prim, tuv

    def trace_noalloc(self, rays, any_hit=False):
        """hj_trace_noalloc: the batch ray query on the library's fixed device scratch (no cudaMalloc; at most
        abi.TRACE_NOALLOC_MAX_RAYS rays per call)."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        n = rays.shape[0]
        prim = np.zeros(n, np.uint32)
        tuv = np.zeros((n, 3), np.float32)
        self._check(self.L.hj_trace_noalloc(self.ctx, rays.ctypes.data_as(abi.f32p), n, int(any_hit), prim.ctypes.data_as(abi.u32p),
                                            tuv.ctypes.data_as(abi.f32p)), "hj_trace_noalloc")
        return prim, tuv

    LIBM_FN = {"sinf": 0, "cosf": 1, "acosf": 2, "logf": 3, "expf": 4, "powf": 5, "atan2f": 6}

    def libm(self, fn, a, b=None):
        """Evaluates one elementary function of csrc/hj_libm.h on the device (hj_libm_eval)."""
        a = np.ascontiguousarray(a, np.float32)
        b = None if b is None else np.ascontiguousarray(b, np.float32)
        out = np.empty_like(a)
        self._check(self.L.hj_libm_eval(self.ctx, self.LIBM_FN[fn], a.ctypes.data_as(abi.f32p), b.ctypes.data_as(abi.f32p) if b is not None else None,
                                        out.ctypes.data_as(abi.f32p), a.size), "hj_libm_eval")
        return out

    # ---- multi-GPU ----
    @staticmethod
    def comm_unique_id():
        L = load_library()
        buf = (C.c_uint8 * 128)()
        if L.hj_comm_unique_id(buf) != 0:
            raise HenjouError("hj_comm_unique_id failed (NCCL unavailable)")
        return bytes(buf)

    def comm_init(self, unique_id, rank, nranks):
        buf = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        self._check(self.L.hj_comm_init(self.ctx, buf, rank, nranks), "hj_comm_init")

    def comm_upload(self, scene, root=0):
        """hj_comm_scene_upload: the geometry crosses PCIe on `root` only and is broadcast to the other ranks over NVLink."""
        self.scene = scene
        desc = scene.to_desc()
        self._check(self.L.hj_comm_scene_upload(self.ctx, C.byref(desc), int(root)), "hj_comm_scene_upload")

    def comm_reduce(self):
        self._check(self.L.hj_comm_reduce(self.ctx), "hj_comm_reduce")

    def close(self):
        if self.ctx:
            self.L.hj_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
