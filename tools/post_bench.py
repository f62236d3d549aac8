#!/usr/bin/env python
"""Times hj_post on a C2 frame (1920x1080) and reports the denoiser's effect: RMSE of 4 spp raw / denoised vs 256 spp."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import henjou_renderer_b200 as hj
from henjou_renderer_b200 import abi

scene, cfg, desc = bench.make_workload(sys.argv[1] if len(sys.argv) > 1 else "c2")
r = hj.Renderer(0)
r.upload(scene)
xf, inv = scene.transforms(cfg["frame"] / 24.0)
r.build_accel(xf, inv)
def frame(spp):
    c = dict(cfg); c["spp"] = spp
    r.render(bench.frame_for(scene, c, spp, 0, spp))
    return r.readback(abi.AOV_COLOR).copy()
ref = frame(256)
noisy = frame(4)
print("non-finite pixels: 256 spp %d, 4 spp %d of %d" % ((~np.isfinite(ref).all(-1)).sum(), (~np.isfinite(noisy).all(-1)).sum(), ref.shape[0] * ref.shape[1]))
good = np.isfinite(ref).all(-1)
def rmse(a):      # relative MSE (Rousselle et al.): HDR fireflies do not drown everything else
    d = np.nan_to_num(a[..., :3], nan=0.0, posinf=1.0)[good] - ref[..., :3][good]
    return float(np.mean(d * d / (ref[..., :3][good] ** 2 + 1e-2)))
for name, mode in (("passthrough", abi.POST_PASSTHROUGH), ("denoise", abi.POST_DENOISE), ("upscale2x", abi.POST_DENOISE_UPSCALE2X)):
    ms = []
    for _ in range(5):
        r.post(mode); ms.append(r.stats()["ms_post"])
    out = r.readback(abi.AOV_OUTPUT)
    print("%-12s ms_post %.3f (min of 5; %d launches)  out %s" % (name, min(ms), r.stats()["post_launches"], out.shape[:2]),
          "" if mode == abi.POST_DENOISE_UPSCALE2X else "relMSE vs 256 spp: raw %.4f -> %.4f" % (rmse(noisy), rmse(out)))
for sc, sn, sa in ((300, 300, 100), (1000, 300, 100), (2000, 300, 100), (4000, 300, 100), (8000, 300, 100), (2000, 100, 50), (2000, 1000, 300), (2000, 300, 30), (2000, 300, 300)):
    r.set_option("denoise_sigma_color_milli", sc); r.set_option("denoise_sigma_normal_milli", sn); r.set_option("denoise_sigma_albedo_milli", sa)
    r.post(abi.POST_DENOISE)
    print("sigma c/n/a %4d %4d %4d: relMSE %.4f" % (sc, sn, sa, rmse(r.readback(abi.AOV_OUTPUT))))
