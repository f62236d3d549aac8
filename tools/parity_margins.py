#!/usr/bin/env python
"""CPU-only parity bridge (TEST INFRASTRUCTURE; runs without a GPU).

The GPU kernels are compared bit for bit against oracle/libhjo.so (the restatement built on the product's shared
hj_libm.h). This tool measures the two remaining links of the chain to the reference's own code:
  (1) oracle/libhjo_glibc.so (the same restatement on glibc's libm)  vs  oracle/_ref/libhjref.so (the reference's own
      headers, glibc): expected bit-equal per path;
  (2) libhjo.so  vs  libhjo_glibc.so: two builds of ONE source that differ only in seven elementary functions, each
      within 1 ulp of the other (tests/test_libm.py). Every per-path difference is therefore caused by <= 1-ulp libm
      differences; the decision traces classify each mismatching path: a discrete decision that flipped (with the drift
      of the compared values between the builds) or the same decisions with an ill-conditioned value.
Writes a markdown table (default profiles/r02_parity_cpu.md).
    python tools/parity_margins.py [--size 192] [--spp 8] [--out profiles/r02_parity_cpu.md]"""
import argparse
import collections
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

import henjou_renderer_b200 as hj  # noqa: E402
from henjou_renderer_b200 import abi, scenes  # noqa: E402
import oracle_binding as ob  # noqa: E402
from oracle_binding import Oracle  # noqa: E402
import parity_tools as pt  # noqa: E402


def cases(size, spp):
    g = os.path.join(ROOT, "tests", "golden")
    c = hj.load_npz(os.path.join(g, "cornelbox.npz")); c.set_sky((0.8, 0.8, 0.8))
    ct = hj.load_npz(os.path.join(g, "cornelbox_texture_test.npz")); ct.set_sky((0.8, 0.8, 0.8))
    cam = dict(position=np.array([4.5, 0.2, 0.1], np.float32), direction=np.array([-1, 0, 0], np.float32), fov=0.48)
    c2 = scenes.rtcamp_like(statue_subdiv=4, detail=0.08)
    c3 = scenes.metaglass_room(subdiv=3)
    c5 = scenes.icosphere_field(n_spheres=300, subdiv=3)
    h = size * 9 // 16
    for name, integ in (("PT", abi.INTEGRATOR_PT), ("NEE", abi.INTEGRATOR_NEE), ("MIS", abi.INTEGRATOR_MIS)):
        yield "C1 cornelbox %s depth 10" % name, c, dict(width=size, height=size, spp=spp, frame=1, max_depth=10, integrator=integ), 1 / 24
    yield "cornelbox_texture_test MIS depth 10", ct, dict(width=size, height=size, spp=spp, frame=0, max_depth=10, camera=cam), 0.0
    yield "C2 stand-in (thin-film, EC-GGX, clearcoat) MIS depth 10", c2, dict(width=size, height=h, spp=spp, frame=1, max_depth=10), 1 / 24
    yield "C3 stand-in (ior-1.5 metaglass) MIS depth 10", c3, dict(width=size, height=h, spp=spp, frame=3, max_depth=10), 0.0
    yield "C5 stand-in (390 k-triangle icosphere field) MIS depth 10", c5, dict(width=size, height=h, spp=max(1, spp // 2), frame=1, max_depth=10), 0.0
    yield "C3 stand-in (ior-1.5 metaglass) MIS depth 32", c3, dict(width=size, height=h, spp=max(1, spp // 2), frame=3, max_depth=32), 0.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=192)
    ap.add_argument("--spp", type=int, default=8)
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r02_parity_cpu.md"))
    args = ap.parse_args()
    have_ref = ob.lib("reference") is not None
    rows, details = [], []
    for name, scene, kw, t in cases(args.size, args.spp):
        xf, inv = scene.transforms(t)
        fparams = hj.frame_params(scene, kw.pop("width"), kw.pop("height"), kw.pop("spp"), **kw)
        o_sh = Oracle(scene, "restatement", xforms=xf, inv_xforms=inv)
        o_gl = Oracle(scene, "glibc", xforms=xf, inv_xforms=inv)
        a = o_sh.render(fparams, debug=True)
        g = o_gl.render(fparams, debug=True)
        npaths = a["path_radiance"][..., 0].size
        ref_col = "n/a (reference hard-codes depth 10)" if fparams.max_depth != 10 else "n/a (oracle/_ref not built)"
        if have_ref and fparams.max_depth == 10:
            r = Oracle(scene, "reference", xforms=xf, inv_xforms=inv).render(fparams, debug=True)
            nbad = int(pt.path_mismatch(g["path_radiance"], r["path_radiance"]).sum())
            ref_col = "bit-equal radiance, same first-hit ids" if pt.bit_equal(g["path_radiance"][..., :3], r["path_radiance"][..., :3]) and pt.bit_equal(g["path_ids"][..., 0], r["path_ids"][..., 0]) \
                else "%d paths > 1e-4 (max rel %.2e)" % (nbad, float(np.nanmax(np.abs(g["path_radiance"][..., :3] - r["path_radiance"][..., :3]) /
                                                                          np.maximum(np.abs(r["path_radiance"][..., :3]), 1e-3))))
        bad = pt.path_mismatch(a["path_radiance"], g["path_radiance"])
        ex = pt.explain_mismatches(o_sh, o_gl, fparams, bad)
        # a "flip" = a threshold decision whose compared values agreed to float rounding in both builds until then; everything else
        # is a continuous quantity that drifted first (an ill-conditioned step), seen later as another triangle or another value
        flips = [e for e in ex if 0 < e["kind"] < 20 and e["first_drift"] is None]
        values = [e for e in ex if not (0 < e["kind"] < 20 and e["first_drift"] is None)]
        drift_after = collections.Counter((e["first_drift"] or {}).get("after", "no recorded quantity drifted > 1e-5") for e in values)
        kinds = collections.Counter(e["name"] for e in flips)
        max_drift = max([e["drift"] for e in flips if e["kind"] < 20], default=0.0)
        rel = np.abs(a["path_radiance"][..., :3].astype(np.float64) - g["path_radiance"][..., :3]) / np.maximum(np.abs(g["path_radiance"][..., :3]), pt.ABS_FLOOR)
        worst_value = max([float(np.nanmax(rel[e["sample"] - fparams.sample_begin, e["pixel"][1], e["pixel"][0]])) for e in values], default=0.0)
        worst_value = min(worst_value, 9.99e9)
        identical = float((a["path_radiance"][..., :3].view(np.uint32) == g["path_radiance"][..., :3].view(np.uint32)).all(-1).mean())
        rows.append((name, npaths, ref_col, identical, int(bad.sum()), bad.mean(), len(flips), "; ".join("%s x%d" % kv for kv in kinds.most_common()) or "-",
                     max_drift, len(values), "; ".join("%s x%d" % kv for kv in drift_after.most_common()) or "-"))
        details += [dict(case=name, **e) for e in ex]
        print(rows[-1], flush=True)
    with open(args.out, "w") as f:
        f.write("# r02 parity bridge on the CPU (tools/parity_margins.py --size %d --spp %d)\n\n" % (args.size, args.spp))
        f.write("Chain: GPU kernels ==(bit-exact, `tests/test_gpu_*`)== `oracle/libhjo.so` (restatement + shared `hj_libm.h`) "
                "--(this table, right half)-- `oracle/libhjo_glibc.so` (same source, glibc libm) ==(this table, left half)== "
                "`oracle/_ref/libhjref.so` (the reference's own `include/kernel/*.h`).\n\n"
                "The two restatement builds differ only in seven elementary functions that agree to <= 1 ulp (`tests/test_libm.py`), so every "
                "per-path difference in the right half is caused by <= 1-ulp libm differences; the decision traces say how: a discrete decision "
                "whose compared values agree to float rounding and straddle the threshold (`drift` = relative distance of the compared values between "
                "the two builds), or a continuous quantity that an ill-conditioned step amplified first (sharp GGX lobes with alpha = 0.01: 1 ulp in a "
                "direction is amplified by ~1/alpha^2; the EC-GGX height walk's pow(1-U, 1/Lambda) with Lambda ~ 0; up to 32 refractions in C3) and that "
                "shows later as another triangle or another value.\n\n")
        f.write("| case | paths | glibc build vs reference headers | shared-libm vs glibc: bit-identical paths | paths > 1e-4 rel | fraction | "
                "flipped threshold decision | which | max drift of the compared values | ill-conditioned continuous step | the drift (> 1e-5 rel) first shows after |\n")
        f.write("|---|---|---|---|---|---|---|---|---|---|---|\n")
        for r in rows:
            f.write("| %s | %d | %s | %.4f | %d | %.2e | %d | %s | %.1e | %d | %s |\n" % r)
        f.write("\nPer-path detail: `%s`.\n" % os.path.basename(args.out).replace(".md", ".json"))
    json.dump(details, open(args.out.replace(".md", ".json"), "w"), indent=0)


if __name__ == "__main__":
    main()
