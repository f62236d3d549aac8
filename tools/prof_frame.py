"""Renders a short C2 frame (few spp) through the C ABI — the command profiled under ncu (profiles/)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import henjou_renderer_b200 as hj

workload = sys.argv[1] if len(sys.argv) > 1 else "c2"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 4
opts = dict(a.split("=") for a in sys.argv[3:])
scene, cfg, desc = bench.make_workload(workload)
cfg["spp"] = spp
r = hj.Renderer(0)
for k, v in opts.items():
    r.set_option(k, int(v))
r.upload(scene)
xf, inv = scene.transforms(cfg["frame"] / 24.0)
r.build_accel(xf, inv)
fp = bench.frame_for(scene, cfg, spp, 0, spp)
r.render(fp)
r.render(fp)
st0 = r.stats()
print("untimed render: ms_render %.2f  Mrays/s %.1f" % (st0["ms_render"], (st0["rays_extend"] + st0["rays_shadow"]) / st0["ms_render"] / 1e3))
if "count_fetches" in opts:
    print({k: v for k, v in r.stats().items() if k in ("nodes_fetched", "tris_fetched", "max_nodes_per_ray", "max_tris_per_ray", "bvh_depth", "bvh_nodes")})
r.set_option("time_stages", 1)
r.render(fp)
st = r.stats()
print(desc, "spp", spp, {k: st[k] for k in ("rays_extend", "rays_shadow", "ms_render", "ms_trace", "ms_shade", "ms_accel", "bvh_nodes")},
      "Mrays/s %.1f" % ((st["rays_extend"] + st["rays_shadow"]) / st["ms_render"] / 1e3))
