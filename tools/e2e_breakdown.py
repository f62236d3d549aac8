#!/usr/bin/env python
"""Where the e2e step of bench.py spends its wall time: upload / accel build / render / read-back, with pageable and
with page-locked (hj_host_pin) host buffers.   python tools/e2e_breakdown.py [workload]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench                                            # noqa: E402
import henjou_renderer_b200 as hj                       # noqa: E402
from henjou_renderer_b200 import abi                    # noqa: E402

scene, cfg, desc = bench.make_workload(sys.argv[1] if len(sys.argv) > 1 else "c2")
W, H = cfg["width"], cfg["height"]
r = hj.Renderer(0)
xf, inv = scene.transforms(cfg["frame"] / 24.0)
fp = bench.frame_for(scene, cfg, cfg["spp"], 0, cfg["spp"])
color = np.empty((H, W, 4), np.float32); rgba = np.empty((H, W, 4), np.uint8)
big = [scene.vertices, scene.normals, scene.texcoords, scene.indices, scene.material_ids, color, rgba]
for mode in ("pageable", "pinned"):
    if mode == "pinned":
        t = time.perf_counter(); r.pin(*big); print("pin: %.1f ms" % (1e3 * (time.perf_counter() - t)))
    acc = np.zeros(4)
    for it in range(4):
        t0 = time.perf_counter(); r.upload(scene); r.sync()
        t1 = time.perf_counter(); r.build_accel(xf, inv); r.sync()
        t2 = time.perf_counter(); r.render(fp)
        t3 = time.perf_counter(); r.readback(abi.AOV_COLOR, color); r.readback(abi.AOV_RGBA8, rgba)
        t4 = time.perf_counter()
        if it:
            acc += np.array([t1 - t0, t2 - t1, t3 - t2, t4 - t3])
    print("%-9s upload %.2f  build %.2f  render %.2f  readback %.2f  total %.2f ms" % ((mode,) + tuple(1e3 * acc / 3) + (1e3 * acc.sum() / 3,)))
r.unpin(*big)
