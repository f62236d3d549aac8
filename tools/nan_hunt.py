#!/usr/bin/env python
"""Finds the non-finite path samples of a frame on the GPU (debug_paths) and re-runs exactly those (pixel row, sample)
through the CPU oracle — the restatement and, where built, the reference's own headers (oracle/_ref) — to tell a kernel
bug from arithmetic the reference shares.   python tools/nan_hunt.py [workload] [spp]      (test tooling: uses oracle/)"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
import henjou_renderer_b200 as hj
from henjou_renderer_b200 import abi
from oracle_binding import Oracle, lib

workload = sys.argv[1] if len(sys.argv) > 1 else "c2"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 64
scene, cfg, desc = bench.make_workload(workload)
W, H = cfg["width"], cfg["height"]
r = hj.Renderer(0)
r.upload(scene)
xf, inv = scene.transforms(cfg["frame"] / 24.0)
r.build_accel(xf, inv)
found = []
chunk = 16
for s0 in range(0, spp, chunk):
    c = dict(cfg); c["spp"] = spp
    fp = bench.frame_for(scene, c, spp, s0, min(spp, s0 + chunk))
    fp.debug_paths = 1
    r.render(fp)
    rad = r.readback(abi.AOV_PATH_RADIANCE)
    bad = np.argwhere(~np.isfinite(rad[..., :3]).all(-1))
    for s, y, x in bad:
        found.append((int(s0 + s), int(y), int(x), [float(v) for v in rad[s, y, x]]))
    del rad
print("non-finite path samples in %d spp of %s: %d" % (spp, workload, len(found)))
out = dict(workload=desc, spp=spp, found=found, oracle=[])
for kind in ("restatement", "reference"):
    if lib(kind) is None:
        print("oracle build '%s' not available" % kind); continue
    o = Oracle(scene, kind=kind, xforms=xf, inv_xforms=inv)
    for s, y, x, g in found[:6]:
        c = dict(cfg); c["spp"] = spp
        fp = bench.frame_for(scene, c, spp, s, s + 1)
        res = o.render(fp, rows=(y, y + 1), debug=True)
        v = res["path_radiance"][0, y, x]
        print("%-11s sample %d pixel (%d,%d): GPU %s  CPU %s" % (kind, s, x, y, g, [float(t) for t in v]))
        out["oracle"].append(dict(kind=kind, sample=s, x=x, y=y, gpu=g, cpu=[float(t) for t in v]))
    o.close()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "nan_hunt_%s.json" % workload), "w"), indent=1, default=str)
