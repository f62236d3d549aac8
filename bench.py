#!/usr/bin/env python
"""bench.py — headline benchmark of the hot path (BASELINE.json): Mrays/s (and spp/s) of the path tracer at
1920x1080 on the rtcamp-like scene, 64 spp, thin-film materials on, MIS, max depth 10 (config C2).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c1|c3|c5]

A step = one whole frame (all pixels x all samples of this rank's slice) through the wavefront integrator.
  value   : whole-job Mrays/s, scene + BVH resident in HBM, device time (CUDA events on the context's stream),
            max over ranks.
  e2e     : the same metric through the C ABI with HOST buffers, every step: hj_scene_upload (geometry H2D) +
            hj_accel_build (transforms H2D, BVH rebuild ≙ updateIASMatrix) + hj_render + hj_readback of the colour
            AOV and the RGBA8 image (D2H), wall clock.
  roofline: traversal kernels (k_trace<0|1|3>): algorithmic bytes = 80 B x nodes fetched + 48 B x triangles fetched
            (SURVEY §8d / BASELINE.md §4; counters-enabled pass) / their launch durations (event pairs, separate pass); the same
            with the 112 B the node really occupies is printed beside it (frac_112B). `traffic` = DRAM bytes per traversal
            launch and the L1/L2 hit rates, issue rate and lanes per instruction come from the committed ncu capture of the
            same workload (profiles/r02_ncu_trace_<workload>.json, written by tools/ncu_summary.py --json).
  cpu_baseline: the reference's own device headers compiled for the host (oracle/_ref, kind "reference") or the
            CPU restatement (kind "port"), all host cores, on a bounded sample (pixel rows x few spp) of the workload.
N > 1 (torchrun): STRONG scaling by default — the fixed 64-spp frame is split into sample slices of 64/N global sample indices
per rank (hj.sample_slice; replicated scene), one NCCL all-reduce of the float4 accumulation buffer per frame inside the timed
region; the line says scaling "strong" at every N and carries the weak-scaling figure (64 spp per rank) as the extra key "weak".
`--scaling weak` times the weak form as the headline instead. `--as-rank-of N` (one GPU) renders what ONE rank of an N-way
strong split renders, to predict the split's efficiency without the box.
`--impl reference` times the CPU reference arm only (rank 0)."""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402


def make_workload(name):
    """-> (scene, dict(width,height,spp,max_depth,integrator,frame), description)"""
    import henjou_renderer_b200 as hj
    from henjou_renderer_b200 import scenes, abi
    if name == "c2":
        s = scenes.rtcamp_like()
        return s, dict(width=1920, height=1080, spp=64, max_depth=10, integrator=abi.INTEGRATOR_MIS, frame=1), \
            "C2 rtcamp-like temple+statue (procedural stand-in, seed 9), 1920x1080, 64 spp, thin-film on, MIS, depth 10"
    if name == "c1":
        s = hj.load_npz(os.path.join(ROOT, "tests", "golden", "cornelbox.npz"))
        s.set_sky((0.8, 0.8, 0.8))
        return s, dict(width=512, height=512, spp=16, max_depth=8, integrator=abi.INTEGRATOR_MIS, frame=1), \
            "C1 cornelbox.gltf 512x512, 16 spp, MIS, depth 8"
    if name == "c3":
        s = scenes.metaglass_room()
        return s, dict(width=1920, height=1080, spp=16, max_depth=32, integrator=abi.INTEGRATOR_MIS, frame=1), \
            "C3 metamaterial-glass room 1920x1080, 16 spp, MIS, depth 32"
    if name == "c5":
        s = scenes.icosphere_field()
        return s, dict(width=3840, height=2160, spp=8, max_depth=10, integrator=abi.INTEGRATOR_MIS, frame=1), \
            "C5 ~10M-triangle icosphere field 3840x2160, 8 spp, MIS, depth 10"
    raise SystemExit("unknown workload " + name)


def frame_for(scene, cfg, spp_total, sample_begin, sample_end):
    import henjou_renderer_b200 as hj
    return hj.frame_params(scene, cfg["width"], cfg["height"], spp_total, frame=cfg["frame"], max_depth=cfg["max_depth"],
                           integrator=cfg["integrator"], sample_begin=sample_begin, sample_end=sample_end)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            pass
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=mx, reasons=sorted(reasons), samples=len(sm))


def cpu_arm(scene, cfg, budget_s=15.0, steps=1, warmup=0):
    """Times the CPU implementation of the path on a bounded sample of the workload: pixel rows around the image
    centre x `spp_s` samples, all host threads. Returns (Mrays/s per step list, info)."""
    import oracle_binding as ob
    kind = "reference" if (ob.lib("reference") is not None and cfg["max_depth"] == 10) else "restatement"
    xf, inv = scene.transforms(cfg["frame"] / 24.0)
    o = ob.Oracle(scene, kind, xforms=xf, inv_xforms=inv)
    W, H = cfg["width"], cfg["height"]
    # explicit thread count: torchrun exports OMP_NUM_THREADS=1, which would leave the CPU arm on one core at N > 1
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    spp_s = min(4, cfg["spp"])
    # size the sample for ~budget_s of CPU work: grow the row band (then the spp) until one render takes >= budget/3
    fp = frame_for(scene, cfg, cfg["spp"], 0, spp_s)
    rows = 8
    while True:
        y0 = (H - rows) // 2
        t = time.time(); o.render(fp, threads=cores, rows=(y0, y0 + rows)); dt = max(time.time() - t, 1e-3)
        if dt >= budget_s / 3 or (rows >= H and spp_s >= cfg["spp"]):
            break
        grow = min(4.0, max(1.5, 0.8 * budget_s / dt))
        if rows < H:
            rows = int(min(H, rows * grow)); rows -= rows % 2
        else:
            spp_s = int(min(cfg["spp"], max(spp_s + 1, spp_s * grow)))
            fp = frame_for(scene, cfg, cfg["spp"], 0, spp_s)
    y0 = (H - rows) // 2
    vals, last = [], None
    for i in range(warmup + steps):
        r = o.render(fp, threads=cores, rows=(y0, y0 + rows))
        rays = r["rays_extend"] + r["rays_shadow"]
        if i >= warmup:
            vals.append(rays / r["seconds"] / 1e6)
        last = r
    info = dict(kind="reference" if kind == "reference" else "port", cores=cores,
                sample="%d pixel rows (y %d..%d) x %d px x %d spp of the workload, %.1f s per step, %d rays" %
                       (rows, y0, y0 + rows, W, spp_s, last["seconds"], last["rays_extend"] + last["rays_shadow"]),
                seconds=last["seconds"], rays=int(last["rays_extend"] + last["rays_shadow"]),
                spp_per_s=rows * W * spp_s / last["seconds"] / (W * H))
    o.close()
    return vals, info


def gpu_arm(args, rank, world, local_rank):
    import henjou_renderer_b200 as hj
    from henjou_renderer_b200 import abi
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    scene, cfg, desc = make_workload(args.workload)
    if args.spp > 0:
        desc = desc.replace("%d spp" % cfg["spp"], "%d spp" % args.spp); cfg["spp"] = args.spp
    W, H, spp = cfg["width"], cfg["height"], cfg["spp"]
    r = hj.Renderer(local_rank)
    for kv in args.opt:                                # kernel-layout options for A/B runs (none changes a result)
        k, v = kv.split("=")
        r.set_option(k, int(v))
    r.upload(scene)
    xf, inv = scene.transforms(cfg["frame"] / 24.0)
    r.build_accel(xf, inv)
    accel_cold_ms = r.stats()["ms_accel"]          # first build of the context: includes the cudaMalloc of every builder buffer
    r.build_accel(xf, inv)
    accel_ms = r.stats()["ms_accel"]               # full rebuild into the buffers the context already owns (the per-frame cost)
    # strong scaling (default): the fixed frame of `spp` samples split into contiguous sample slices, one per rank;
    # weak scaling: every rank renders `spp` samples of a frame with spp*world global sample indices
    strong = args.scaling != "weak"
    fp_weak = frame_for(scene, cfg, *hj.weak_slice(spp, rank, world))
    fp_strong = frame_for(scene, cfg, spp, *hj.sample_slice(spp, rank, world))
    if args.as_rank_of > 1:
        fp_strong = frame_for(scene, cfg, spp, *hj.sample_slice(spp, 0, args.as_rank_of))
    fp = fp_strong if strong else fp_weak
    if world > 1:
        import torch
        uid = [hj.Renderer.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        r.comm_init(uid[0], rank, world)

    def barrier():
        if world > 1:
            import torch
            dist.barrier()
            torch.cuda.synchronize()

    def step(fp):
        r.render(fp, sync=False)
        if world > 1:
            r.comm_reduce()          # NCCL all-reduce of the accumulation buffer + finalize (syncs the stream)
        r.sync()

    step_wall = []

    def timed(fp, steps, warmup):
        """-> (seconds: max over ranks, rays: sum over ranks, launches: sum over ranks, wall, per-rank device ms list)"""
        for _ in range(warmup):
            step(fp)
        barrier()
        t0 = time.perf_counter()
        dev_ms, rays, launches = 0.0, 0, 0
        step_wall.clear()
        for _ in range(steps):
            ts = time.perf_counter()
            step(fp)
            step_wall.append(round(1e3 * (time.perf_counter() - ts), 3))
            st = r.stats()
            dev_ms += st["ms_render"]; rays += st["rays_extend"] + st["rays_shadow"]; launches += st["kernel_launches"]
        barrier()
        wall = time.perf_counter() - t0
        # N = 1: CUDA events on the context stream. N > 1: barrier-bracketed wall clock (render + all-reduce + finalize), max over ranks
        t_rank = wall if world > 1 else dev_ms / 1e3
        if world > 1:
            import torch
            tt = torch.tensor([t_rank, float(rays), float(launches), dev_ms], dtype=torch.float64, device="cuda")
            tmax = tt.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            tmin = tt.clone(); dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
            tsum = tt.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
            return float(tmax[0]), int(tsum[1]), int(tsum[2]), wall, (float(tmin[3]) / steps, float(tmax[3]) / steps)
        return t_rank, rays, launches, wall, (dev_ms / steps, dev_ms / steps)

    # only rank 0's record is reported: one nvidia-smi poller per box, not one per rank (its queries perturb the driver; r02: at N = 2 / 4
    # single steps of 258 instead of 175 ms / 100 instead of 89 ms with a poller on every rank)
    clocks = ClockSampler(local_rank) if rank == 0 else None
    t_all, rays_all, launches_all, wall, render_ms = timed(fp, args.steps, args.warmup)
    clk = clocks.stop() if clocks else None
    value = rays_all / t_all / 1e6
    out = dict(metric="Mrays/s", value=round(value, 2), unit="Mrays/s", n_gpus=world, steps=args.steps, warmup=args.warmup,
               ms_per_step=round(1e3 * t_all / args.steps, 3), higher_is_better=True, scaling="strong" if strong else "weak", vs_baseline=None, dtype="f32",
               data="synthetic", impl="ours",
               config=dict(workload=desc, triangles=scene.n_triangles, width=W, height=H, spp_per_gpu=(fp.sample_end - fp.sample_begin), spp_frame=int(fp.spp),
                           max_depth=cfg["max_depth"],
                           integrator="MIS", parallelism=("sample slices of the fixed %d-spp frame, " % spp if strong else "%d spp per rank, " % spp) +
                           "replicated scene, NCCL all-reduce of accum" if world > 1 else ("single GPU" if args.as_rank_of <= 1 else
                                                                                        "single GPU rendering rank 0's slice of a %d-way strong split" % args.as_rank_of),
                           l2_policy="inputs larger than L2 (126 MB): %d passes of %.1f M paths in flight x ~280 B of path state rewritten every "
                                     "bounce + %.0f MB of BVH nodes and triangles" % (r.stats()["passes"], (fp.sample_end - fp.sample_begin) * W * H / max(1, r.stats()["passes"]) / 1e6,
                                                                                     (r.stats()["bvh_nodes"] * r.stats()["bvh_node_bytes"] + scene.n_triangles * 48) / 1e6),
                           timing="CUDA events on the context stream (N=1) / barrier-bracketed wall clock incl. all-reduce, max over ranks (N>1)"),
               spp_per_s=round(int(fp.spp) * args.steps / t_all, 3) if args.as_rank_of <= 1 else None, rays_per_step=rays_all // args.steps,
               render_ms_per_rank_min_max=[round(x, 3) for x in render_ms], step_wall_ms_rank0=list(step_wall),
               gpu_launches=launches_all, accel_build_ms=round(accel_ms, 3), accel_first_build_ms=round(accel_cold_ms, 3), clocks=clk, wall_s=round(wall, 3))
    if world > 1 and strong and not args.no_weak:
        # the weak-scaling figure beside it: every rank renders the full `spp` samples of a spp*N-sample frame
        tw, rw, _, _, _ = timed(fp_weak, args.steps, 1)
        out["weak"] = dict(value=round(rw / tw / 1e6, 2), unit="Mrays/s", ms_per_step=round(1e3 * tw / args.steps, 3), spp_per_gpu=spp, spp_frame=spp * world)
    # ---- e2e through the C ABI with host buffers, every rank: upload + BVH rebuild + render (+ all-reduce) + read-back
    h2d = scene.vertices.nbytes + scene.normals.nbytes + scene.texcoords.nbytes + scene.indices.nbytes + scene.material_ids.nbytes + \
        scene.prim_offsets.nbytes + xf.nbytes + inv.nbytes + (scene.lut.nbytes if scene.lut is not None else 0) + scene.ibl.nbytes
    d2h = W * H * 16 + W * H * 4
    color = np.empty((H, W, 4), np.float32); rgba = np.empty((H, W, 4), np.uint8)
    # the host buffers are page-locked once (hj_host_pin), as a host that re-uploads every frame would do
    pinned = [a for a in (scene.vertices, scene.normals, scene.texcoords, scene.indices, scene.material_ids, color, rgba)
              if a.flags["C_CONTIGUOUS"]] if not args.pageable else []
    r.pin(*pinned)
    def e2e_step():
        if world > 1:
            r.comm_upload(scene, 0)          # hj_comm_scene_upload: one PCIe copy of the geometry on rank 0, ncclBroadcast over NVLink to the others
        else:
            r.upload(scene)
        r.build_accel(xf, inv); r.render(fp, sync=False)
        if world > 1:
            r.comm_reduce()
        r.sync()
        r.readback(abi.AOV_COLOR, color); r.readback(abi.AOV_RGBA8, rgba)
        st = r.stats()
        return st["rays_extend"] + st["rays_shadow"]

    for _ in range(args.e2e_warmup):         # untimed: NCCL sets up the broadcast's connections on first use (8 ranks: ~1 s)
        e2e_step()
    barrier()
    e_rays, t0, e_steps = 0, time.perf_counter(), []
    for _ in range(args.steps):
        ts = time.perf_counter()
        e_rays += e2e_step()
        e_steps.append(round(1e3 * (time.perf_counter() - ts), 2))
    barrier()
    e_t = time.perf_counter() - t0
    if world > 1:
        import torch
        tt = torch.tensor([e_t, float(e_rays)], dtype=torch.float64, device="cuda")
        tmax = tt.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = tt.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        e_t, e_rays = float(tmax[0]), int(tsum[1])
    small = scene.prim_offsets.nbytes + xf.nbytes + inv.nbytes + (scene.lut.nbytes if scene.lut is not None else 0) + scene.ibl.nbytes
    out["e2e"] = dict(value=round(e_rays / e_t / 1e6, 2), unit="Mrays/s", h2d_bytes_per_step=int(h2d) + int(small) * (world - 1), d2h_bytes_per_step=int(d2h) * world,
                      nvlink_broadcast_bytes_per_step=int(h2d - small) * (world - 1) if world > 1 else 0,
                      ms_per_step=round(1e3 * e_t / args.steps, 3), ms_steps_rank0=e_steps, checksum=float(np.nanmean(np.where(np.isfinite(color[..., :3]), color[..., :3], np.nan))),
                      non_finite_pixels=int((~np.isfinite(color[..., :3]).all(-1)).sum()),
                      host_memory="pageable" if args.pageable else "pinned (hj_host_pin)")
    r.unpin(*pinned)
    if rank == 0 and world == 1 and not args.no_roofline:
        # ---- roofline of the traversal kernels: bytes from a counters pass, time from an event-pair pass
        r.set_option("count_fetches", 1); r.render(fp); sc = r.stats(); r.set_option("count_fetches", 0)
        r.set_option("time_stages", 1); r.render(fp); stt = r.stats(); r.set_option("time_stages", 0)
        # SURVEY §8d / BASELINE.md §4: B_ray = 80 N_node + 48 N_tri (the 80-B node that was planned). The node built here is 112 B
        # (the eight child masks moved into it); that figure is printed beside the headline as frac_112B.
        algo = 80.0 * sc["nodes_fetched"] + 48.0 * sc["tris_fetched"]
        algo112 = float(sc["bvh_node_bytes"]) * sc["nodes_fetched"] + 48.0 * sc["tris_fetched"]
        nrays = sc["rays_extend"] + sc["rays_shadow"]
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        # The timed region runs 8 wavefront pipelines concurrently, so a traversal launch has no private duration there.
        # Its share of the GPU is taken from the event-pair pass, where every launch runs alone on one stream
        # (share = sum of traversal launch durations / sum of all stage durations), and applied to the timed step.
        share = stt["ms_trace"] / max(stt["ms_trace"] + stt["ms_shade"], 1e-9)
        step_ms = 1e3 * t_all / args.steps
        n_types = {abi.INTEGRATOR_PT: 1, abi.INTEGRATOR_NEE: 2, abi.INTEGRATOR_MIS: 3}[cfg["integrator"]]
        n_trace_launches = n_types * cfg["max_depth"] * max(1, stt["passes"])
        ach = algo / (share * step_ms / 1e3) / 1e9
        ach_serial = algo / (stt["ms_trace"] / 1e3) / 1e9
        ach112 = algo112 / (share * step_ms / 1e3) / 1e9
        # measured DRAM traffic and the counters that say what bounds the kernel: from the committed ncu capture of this workload
        ncu, ncu_path = {}, os.path.join(ROOT, "profiles", "r02_ncu_trace_%s.json" % args.workload)
        try:
            ncu = json.load(open(ncu_path))
        except Exception:
            pass
        tr = ncu.get("traversal", {})
        out["roofline"] = dict(bound="hbm", achieved=round(ach, 1), peak=peak, unit="GB/s", frac=round(ach / peak, 4),
                               traffic=tr.get("dram_bytes_per_launch"), traffic_source=os.path.relpath(ncu_path, ROOT) if tr else None,
                               traffic_note=tr.get("note"),
                               frac_112B=round(ach112 / peak, 4), achieved_112B=round(ach112, 1),
                               dram_gbs_ncu=tr.get("dram_gbs"), dram_frac_of_peak_ncu=round(tr["dram_gbs"] / peak, 4) if tr.get("dram_gbs") else None,
                               algorithmic_bytes_per_launch_ncu=tr.get("algorithmic_bytes_per_launch"),
                               l1_hit=tr.get("l1_hit"), l2_hit=tr.get("l2_hit"), issue_active=tr.get("issue_active"), lanes_per_inst=tr.get("lanes_per_inst"),
                               bound_by=tr.get("bound_by"),
                               peak_source="MEASURED_PEAKS.json" if "hbm_gbs" in peaks else "fallback",
                               kernel="k_trace<0|1|3> (BVH traversal: extend, shadow, MIS emitter rays)",
                               byte_definition="80 B x nodes + 48 B x triangles fetched (SURVEY 8d); frac_112B uses the 112 B a node occupies",
                               method="algorithmic bytes per step / (traversal share x timed step); share from the serial event-pair pass",
                               trace_share=round(share, 4), achieved_serial_launches=round(ach_serial, 1),
                               serial_trace_ms_per_step=round(stt["ms_trace"], 3), serial_shade_ms_per_step=round(stt["ms_shade"], 3),
                               serial_step_ms=round(stt["ms_render"], 3), trace_launches_per_step=n_trace_launches,
                               avg_launch_us_serial=round(1e3 * stt["ms_trace"] / n_trace_launches, 1),
                               nodes_per_ray=round(sc["nodes_fetched"] / nrays, 2), tris_per_ray=round(sc["tris_fetched"] / nrays, 2),
                               bytes_per_ray=round(algo / nrays, 1), bytes_per_ray_112B=round(algo112 / nrays, 1), bound_mrays=round(peak * 1e9 / (algo / nrays) / 1e6, 1),
                               max_nodes_one_ray=sc["max_nodes_per_ray"], max_tris_one_ray=sc["max_tris_per_ray"], bvh_depth=sc["bvh_depth"])
        if not args.no_cpu:
            vals, info = cpu_arm(scene, cfg, budget_s=args.cpu_budget)
            out["cpu_baseline"] = dict(value=round(vals[0], 3), unit="Mrays/s", cores=info["cores"], kind=info["kind"], sample=info["sample"],
                                       spp_per_s=round(info["spp_per_s"], 5))
    if rank == 0:
        print(json.dumps(out), flush=True)
    r.close()
    if world > 1:
        dist.destroy_process_group()


def reference_arm(args, rank, world):
    if rank != 0:
        return
    scene, cfg, desc = make_workload(args.workload)
    if args.spp > 0:
        desc = desc.replace("%d spp" % cfg["spp"], "%d spp" % args.spp); cfg["spp"] = args.spp
    vals, info = cpu_arm(scene, cfg, budget_s=args.cpu_budget, steps=args.steps, warmup=min(args.warmup, 1))
    v = float(np.mean(vals))
    out = dict(metric="Mrays/s", value=round(v, 3), unit="Mrays/s", n_gpus=world, steps=args.steps, warmup=args.warmup,
               ms_per_step=round(1e3 * info["seconds"], 3), higher_is_better=True, scaling=args.scaling, vs_baseline=None, dtype="f32",
               data="synthetic", impl="reference",
               config=dict(workload=desc, triangles=scene.n_triangles, width=cfg["width"], height=cfg["height"], max_depth=cfg["max_depth"], integrator="MIS"),
               cpu_baseline=dict(value=round(v, 3), unit="Mrays/s", cores=info["cores"], kind=info["kind"], sample=info["sample"]),
               e2e=dict(value=round(v, 3), unit="Mrays/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0,
               note="reference's own device headers compiled for the host (oracle/_ref) with the reconstructed raygen/closest-hit and a CPU BVH; "
                    "the reference through OptiX/RT cores is unavailable on this box")
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--cpu-budget", type=float, default=12.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--pageable", action="store_true", help="e2e leg with pageable host buffers (no hj_host_pin)")
    ap.add_argument("--opt", action="append", default=[], help="hj_set_option key=value (repeatable), for A/B runs")
    ap.add_argument("--spp", type=int, default=0, help="override the workload's samples per pixel (C5 as BASELINE.json states it: --workload c5 --spp 1024)")
    ap.add_argument("--e2e-warmup", type=int, default=1, help="untimed end-to-end steps before the timed ones")
    ap.add_argument("--no-weak", action="store_true", help="N > 1: skip the extra weak-scaling measurement (long workloads)")
    ap.add_argument("--no-roofline", action="store_true", help="skip the counters / stage-timing passes and the CPU baseline (long workloads)")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"], help="N > 1: split the fixed frame (strong, default) or 64 spp per rank (weak)")
    ap.add_argument("--as-rank-of", type=int, default=1, help="one GPU: render rank 0's slice of an N-way strong split (predicts the split without the box)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        reference_arm(args, rank, world)
    else:
        gpu_arm(args, rank, world, local)


if __name__ == "__main__":
    main()
