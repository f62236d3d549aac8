"""Camera and per-frame launch parameters: the host half of the reference's frame loop
(/root/reference/include/renderer/renderer.h:1126-1227) restated with float32 arithmetic."""
import math
import numpy as np

from . import abi

f32 = np.float32


def _cross(a, b):
    return np.array([a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]], f32)


def camera_basis(position, direction, fov, animation=None, time=0.0):
    """Returns (pos, dir, up, right, camera_f). fov in radians; camera_f = 2/tan(fov) — fov, NOT fov/2
    (renderer.h:1147). Animated branch renderer.h:1149-1162, static branch :1163-1169 (un-normalised)."""
    camera_f = f32(2.0 / math.tan(float(f32(fov))))
    position = np.asarray(position, f32)
    direction = np.asarray(direction, f32)
    if animation is not None:
        A = animation.affine(time)
        R = animation.rotate_only(time)

        def mul(M, v4):   # common/matrix.h operator*(Affine4x4, float4): left-to-right float accumulation
            out = np.zeros(4, f32)
            for r in range(4):
                acc = f32(v4[0] * M[r, 0])
                for k in (1, 2, 3):
                    acc = f32(acc + f32(v4[k] * M[r, k]))
                out[r] = acc
            return out

        pos = mul(A, np.array([position[0], position[1], position[2], 1.0], f32))[:3]
        d = mul(R, np.array([direction[0], direction[1], direction[2], 0.0], f32))[:3]
        up = mul(R, np.array([0.0, 1.0, 0.0, 0.0], f32))[:3]
        right = _cross(d, up)
        inv = f32(1.0) / f32(np.sqrt(f32(right[0] * right[0] + right[1] * right[1] + right[2] * right[2])))
        right = (right * inv).astype(f32)
        return pos, d, up, right, camera_f
    right = _cross(direction, np.array([0, 1, 0], f32))
    up = _cross(right, direction)
    return position, direction, up, right, camera_f


def frame_params(scene, width, height, spp, frame=0, fps=24, integrator=abi.INTEGRATOR_MIS, max_depth=10,
                 tonemap=abi.TONEMAP_NONE, pixel_jitter=1, debug_paths=0, ibl_intensity=1.0,
                 sample_begin=0, sample_end=None, allow_camera_animation=True, camera=None, sky_sampling=0):
    """hj_frame_params for frame index `frame` (time = frame/fps, renderer.h:1128; scramble = frame)."""
    cam = camera if camera is not None else scene.camera
    time = float(f32(frame) / f32(fps))
    anim = cam.get("animation") if allow_camera_animation else None
    pos, d, up, right, cf = camera_basis(cam["position"], cam["direction"], cam["fov"], anim, time)
    fp = abi.hj_frame_params()
    fp.width, fp.height, fp.spp = width, height, spp
    fp.sample_begin = sample_begin
    fp.sample_end = spp if sample_end is None else sample_end
    fp.frame = frame
    for name, v in (("camera_pos", pos), ("camera_dir", d), ("camera_up", up), ("camera_right", right)):
        setattr(fp, name, (abi.C.c_float * 3)(float(v[0]), float(v[1]), float(v[2])))
    fp.camera_f = float(cf)
    fp.ibl_intensity = ibl_intensity
    fp.integrator, fp.max_depth, fp.tonemap = integrator, max_depth, tonemap
    fp.pixel_jitter, fp.debug_paths = pixel_jitter, debug_paths
    fp.sky_sampling = sky_sampling        # 0 = reference behaviour; 1 = light-sampled sky with MIS (extension, MIS integrator only)
    return fp
