"""Renders a short frame of a bench workload through the C ABI — the command profiled under ncu (profiles/).

    python tools/prof_frame.py <workload c1|c2|c3|c5> <spp> [option=value ...] [--one-frame] [--stats-json out.json]
  default: two untimed frames, then a time_stages frame (serial event pairs) and its Mrays/s;
  --one-frame: exactly one frame after the accel build (the run captured by `ncu --set full -k regex:k_trace|k_surface|k_nee|k_sample`);
  --stats-json: a count_fetches frame first; writes nodes / triangles fetched, traversal launches and the algorithmic bytes per
                traversal launch (80 B x nodes + 48 B x triangles) of THIS frame, for tools/ncu_summary.py --algo-bytes-per-launch."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import henjou_renderer_b200 as hj
from henjou_renderer_b200 import abi

argv = sys.argv[1:]
stats_json = None
if "--stats-json" in argv:
    k = argv.index("--stats-json")
    stats_json = argv[k + 1]
    del argv[k:k + 2]
flags = [a for a in argv if a.startswith("--")]
args = [a for a in argv if not a.startswith("--")]
workload = args[0] if args else "c2"
spp = int(args[1]) if len(args) > 1 else 4
opts = dict(a.split("=") for a in args[2:])
scene, cfg, desc = bench.make_workload(workload)
cfg["spp"] = spp
r = hj.Renderer(0)
for k, v in opts.items():
    r.set_option(k, int(v))
r.upload(scene)
xf, inv = scene.transforms(cfg["frame"] / 24.0)
r.build_accel(xf, inv)
fp = bench.frame_for(scene, cfg, spp, 0, spp)
if "--one-frame" in flags:
    r.render(fp)
    st = r.stats()
    print(desc, "spp", spp, "one frame: ms_render %.2f  Mrays/s %.1f  launches %d" % (st["ms_render"], (st["rays_extend"] + st["rays_shadow"]) / st["ms_render"] / 1e3, st["kernel_launches"]))
    sys.exit(0)
if stats_json:
    r.set_option("count_fetches", 1); r.render(fp); sc = r.stats(); r.set_option("count_fetches", 0)
    n_types = {abi.INTEGRATOR_PT: 1, abi.INTEGRATOR_NEE: 2, abi.INTEGRATOR_MIS: 3}[cfg["integrator"]]
    launches = n_types * cfg["max_depth"] * max(1, sc["passes"])
    algo = 80.0 * sc["nodes_fetched"] + 48.0 * sc["tris_fetched"]
    json.dump(dict(workload=desc, spp=spp, options=opts, nodes_fetched=sc["nodes_fetched"], tris_fetched=sc["tris_fetched"], rays=sc["rays_extend"] + sc["rays_shadow"],
                   traversal_launches=launches, algorithmic_bytes_80B=algo, algorithmic_bytes_per_launch_80B=algo / launches,
                   algorithmic_bytes_per_launch_112B=(112.0 * sc["nodes_fetched"] + 48.0 * sc["tris_fetched"]) / launches,
                   bvh_nodes=sc["bvh_nodes"], bvh_bytes=sc["bvh_nodes"] * sc["bvh_node_bytes"] + 48 * scene.n_triangles,
                   max_nodes_per_ray=sc["max_nodes_per_ray"], max_tris_per_ray=sc["max_tris_per_ray"]), open(stats_json, "w"), indent=1)
    print(open(stats_json).read())
r.render(fp)
r.render(fp)
st0 = r.stats()
print("untimed render: ms_render %.2f  Mrays/s %.1f" % (st0["ms_render"], (st0["rays_extend"] + st0["rays_shadow"]) / st0["ms_render"] / 1e3))
r.set_option("time_stages", 1)
r.render(fp)
st = r.stats()
print(desc, "spp", spp, {k: st[k] for k in ("rays_extend", "rays_shadow", "ms_render", "ms_trace", "ms_shade", "ms_accel", "bvh_nodes")},
      "Mrays/s %.1f" % ((st["rays_extend"] + st["rays_shadow"]) / st["ms_render"] / 1e3))
