"""C4 (BASELINE.json configs[3]): animated rtcamp-like sequence at 1920x1080, per-frame accel update + render + RGBA8
read-back through the C ABI, against the fps.txt time budget of the reference's frame loop (renderer.h:1126-1310,
the budget check it comments out at :1247-1254).  Compares the two per-frame update policies of hj_accel_build
(full rebuild vs. refit of the first frame's topology) and finds the spp that fits the budget.

    python tools/sequence_bench.py [frames=24] [fps=24] [spp=auto|N] [animated=24] [bands=N] [pipelines=N]  -> one JSON line
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import henjou_renderer_b200 as hj  # noqa: E402
from henjou_renderer_b200 import abi, scenes  # noqa: E402

opts = dict(a.split("=") for a in sys.argv[1:])
frames = int(opts.get("frames", 24))
fps = float(opts.get("fps", 24))
animated = int(opts.get("animated", 24))
W, H = int(opts.get("width", 1920)), int(opts.get("height", 1080))
budget_ms = 1000.0 / fps

scene = scenes.rtcamp_like(animated=animated)
r = hj.Renderer(0)
for k in ("bands", "pipelines", "leaf_split", "builder"):      # kernel options for A/B runs
    if k in opts:
        r.set_option(k, int(opts[k]))
r.upload(scene)
rgba = np.empty((H, W, 4), np.uint8)
r.pin(rgba)


def run_sequence(spp, policy):
    """-> per-frame dicts (wall ms of update+render+read-back, device ms of the accel update and the render)"""
    out = []
    overlapped = policy.endswith("_async")                 # next frame's update enqueued under this frame's shading (hj_accel_build_async)
    if overlapped:
        r.build_accel(*scene.transforms(0.0), mode=abi.ACCEL_BUILD)
    for f in range(frames):
        fp = hj.frame_params(scene, W, H, spp, frame=f, max_depth=10)
        t0 = time.perf_counter()
        if overlapped:
            r.render(fp, sync=False)
            if f + 1 < frames:
                r.build_accel(*scene.transforms((f + 1) / fps), mode=abi.ACCEL_BUILD if policy == "rebuild_async" else abi.ACCEL_REFIT, sync=False)
            r.sync()
        else:
            xf, inv = scene.transforms(f / fps)
            mode = abi.ACCEL_BUILD if (policy == "rebuild" or f == 0) else abi.ACCEL_REFIT
            r.build_accel(xf, inv, mode=mode)
            r.render(fp)
        r.readback(abi.AOV_RGBA8, out=rgba)
        wall = (time.perf_counter() - t0) * 1e3
        st = r.stats()
        out.append(dict(frame=f, wall_ms=wall, accel_ms=float(st["ms_accel"]), render_ms=float(st["ms_render"]),
                        rays=int(st["rays_extend"]) + int(st["rays_shadow"]), checksum=int(rgba[::7, ::11].astype(np.uint64).sum())))
    return out


def summary(rows):
    body = rows[1:] if len(rows) > 1 else rows          # frame 0 always builds
    wall = np.array([x["wall_ms"] for x in body])
    return dict(frames=len(rows), wall_ms_mean=round(float(wall.mean()), 3), wall_ms_max=round(float(wall.max()), 3),
                accel_ms_mean=round(float(np.mean([x["accel_ms"] for x in body])), 3),
                render_ms_mean=round(float(np.mean([x["render_ms"] for x in body])), 3),
                mrays_per_s=round(float(sum(x["rays"] for x in body) / wall.sum() / 1e3), 1),
                frames_over_budget=int((wall > budget_ms).sum()))


run_sequence(1, "rebuild")                               # warm-up (allocations, NCCL-free)
if opts.get("spp", "auto") == "auto":
    # largest spp whose slowest frame fits the budget: time 2 spp, extrapolate, verify, step down if needed
    probe = summary(run_sequence(2, "refit_async"))
    per_spp = max(probe["render_ms_mean"] / 2.0, 1e-3)
    spp = max(1, int((budget_ms - probe["accel_ms_mean"] - 2.0) / per_spp))
    fits = lambda n: summary(run_sequence(n, "refit_async"))["frames_over_budget"] == 0
    while spp > 1 and not fits(spp):
        spp -= 1
    while spp < 64 and fits(spp + 1):                    # the 2-spp probe overestimates the cost per sample: try upwards too
        spp += 1
else:
    spp = int(opts["spp"])

res = {}
checks = {}
for policy in ("rebuild", "refit", "rebuild_async", "refit_async"):
    rows = run_sequence(spp, policy)
    res[policy] = summary(rows)
    checks[policy] = [x["checksum"] for x in rows]
# refit keeps frame 0's topology: traversal order differs, the closest-hit rule does not -> identical frames
res["frames_identical_rebuild_vs_refit"] = checks["rebuild"] == checks["refit"]
res["frames_identical_blocking_vs_async"] = checks["rebuild"] == checks["rebuild_async"] == checks["refit_async"]
line = dict(workload="C4 rtcamp-like stand-in, %d rigidly animated instances, %d frames at %g fps, %dx%d, MIS, depth 10"
            % (animated, frames, fps, W, H), triangles=scene.n_triangles, budget_ms=round(budget_ms, 3), spp=spp, **res)
print(json.dumps(line))
r.unpin(rgba)
r.close()
