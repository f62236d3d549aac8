"""Shared by tests/ and tools/parity_margins.py (TEST INFRASTRUCTURE): per-path comparison of two renders and the
alignment of two decision traces (oracle_binding.Oracle.path_events) of one path."""
import numpy as np

import oracle_binding as ob

REL = 1e-4          # north_star: per-path radiance within 1e-4 relative
ABS_FLOOR = 1e-3    # radiance below this is compared absolutely at REL*ABS_FLOOR


def path_mismatch(got, want):
    """bool[...]: paths whose radiance differs by more than REL relative (NaN == NaN counts as equal)."""
    a, b = got[..., :3].astype(np.float64), want[..., :3].astype(np.float64)
    with np.errstate(invalid="ignore"):
        bad = np.abs(a - b) > REL * np.maximum(np.abs(b), ABS_FLOOR)
    bad |= np.isnan(a) != np.isnan(b)
    return bad.any(-1)


def bit_equal(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint8), np.ascontiguousarray(b).view(np.uint8))


def log_measurement(**kw):
    """Appends one JSON line to gpurun_out/r02_parity_gpu.jsonl (the measured mismatch fractions that profiles/r02_parity.md
    quotes come from here)."""
    import json, os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    path = os.environ.get("HJ_PARITY_LOG") or os.path.join(root, "gpurun_out", "r02_parity_gpu.jsonl")
    try:
        os.makedirs(os.path.dirname(path), exist_ok=True)
        with open(path, "a") as f:
            f.write(json.dumps({k: (v.item() if hasattr(v, "item") else v) for k, v in kw.items()}) + "\n")
    except OSError:
        pass


def compare_paths(label, got_rad, got_ids, want, exact=True, bound=0.0):
    """Per-path comparison of a GPU frame (HJ_AOV_PATH_RADIANCE / HJ_AOV_PATH_IDS) with an oracle frame of the same rows.
    exact=True (oracle/libhjo.so, same libm as the kernels): radiance rgb, the per-path ray count and the hash of every
    closest-hit id must be the same bits. exact=False (oracle/_ref or the glibc build): first-hit ids equal, and at most
    `bound` of the paths beyond 1e-4 relative (the libm-difference fraction measured in profiles/r02_parity_cpu.md)."""
    wr, wi = want["path_radiance"], want["path_ids"]
    assert got_rad.shape == wr.shape and got_ids.shape == wi.shape
    first_ok = np.array_equal(got_ids[..., 0], wi[..., 0])
    rgb_same = ((got_rad[..., :3].view(np.uint32) == wr[..., :3].view(np.uint32)) | (np.isnan(got_rad[..., :3]) & np.isnan(wr[..., :3]))).all(-1)
    bad = path_mismatch(got_rad, wr)
    n = rgb_same.size
    rec = dict(case=label, exact=exact, paths=int(n), first_hit_ids_equal=bool(first_ok), radiance_bit_identical=float(rgb_same.mean()),
               paths_beyond_1e4=int(bad.sum()), fraction_beyond_1e4=float(bad.mean()), bound=bound)
    if exact:
        rec["hash_equal"] = float((got_ids[..., 1] == wi[..., 1]).mean())
        rec["ray_count_equal"] = float((got_rad[..., 3] == wr[..., 3]).mean())
    log_measurement(**rec)
    assert first_ok, "%s: first-hit primitive ids differ" % label
    if exact:
        assert rgb_same.all(), "%s: %d of %d paths are not bit-identical (%d beyond 1e-4 relative)" % (label, int((~rgb_same).sum()), n, int(bad.sum()))
        assert np.array_equal(got_ids[..., 1], wi[..., 1]), "%s: closest-hit id hashes differ" % label
        assert np.array_equal(got_rad[..., 3], wr[..., 3]), "%s: per-path ray counts differ" % label
    else:
        assert bad.mean() <= bound, "%s: %.3e of the paths beyond 1e-4 relative (bound %.1e)" % (label, bad.mean(), bound)
    return bad


def _rel(x, y):
    m = max(abs(float(x)), abs(float(y)))
    return 0.0 if m == 0 else abs(float(x) - float(y)) / m


def first_divergence(oracle_a, oracle_b, fparams, pixel, sample):
    """Aligns the decision traces of one path in two oracle builds. Returns a dict describing the first event whose
    outcome differs: kind, name, index, the compared values in both builds, `drift` = how far the two builds' values of
    the compared quantities are apart (relative) and `margin` = how close the decision was to its threshold in build a
    (relative). For a closest-hit event: the smallest barycentric of the hit (distance to the triangle's edge) in both."""
    ea, ra = oracle_a.path_events(fparams, pixel, sample)
    eb, rb = oracle_b.path_events(fparams, pixel, sample)
    n = min(len(ea), len(eb))
    # where the continuous state of the two builds first drifts apart by more than float rounding (1e-5 relative), and
    # after which kind of event: an ill-conditioned step (a sharp GGX lobe, pow(x, 1/Lambda) with Lambda ~ 0) shows here
    first_drift = None
    for i in range(n):
        x, y = ea[i], eb[i]
        if first_drift is None and x["kind"] == y["kind"] and x["kind"] != 13 and max(_rel(x["a"], y["a"]), _rel(x["b"], y["b"])) > 1e-5:
            first_drift = dict(index=i, name=ob.EVENT_NAMES.get(int(x["kind"]), "?"), after=ob.EVENT_NAMES.get(int(ea[i - 1]["kind"]), "?") if i else "camera ray",
                               rel=max(_rel(x["a"], y["a"]), _rel(x["b"], y["b"])))
    for i in range(n):
        x, y = ea[i], eb[i]
        if x["kind"] != y["kind"] or x["outcome"] != y["outcome"]:
            k = int(x["kind"])
            d = dict(index=i, kind=k, name=ob.EVENT_NAMES.get(k, str(k)), same_kind=bool(x["kind"] == y["kind"]),
                     a=(float(x["a"]), float(x["b"])), b=(float(y["a"]), float(y["b"])), outcome=(int(x["outcome"]), int(y["outcome"])),
                     radiance=(ra.tolist(), rb.tolist()), first_drift=first_drift if first_drift and first_drift["index"] < i else None)
            if k < 20:
                d["drift"] = max(_rel(x["a"], y["a"]), _rel(x["b"], y["b"]))
                d["margin"] = _rel(x["a"], x["b"])
            else:
                d["drift"] = _rel(x["b"], y["b"])                   # hit distance
                d["margin"] = min(abs(float(x["a"])), abs(float(y["a"])))   # smallest barycentric: distance to an edge
            return d
    return dict(index=n, kind=0, name="no diverging decision (values only)" if len(ea) == len(eb) else "trace length", same_kind=True,
                a=(0.0, 0.0), b=(0.0, 0.0), outcome=(len(ea), len(eb)), drift=0.0, margin=0.0, radiance=(ra.tolist(), rb.tolist()), first_drift=first_drift)


def explain_mismatches(oracle_a, oracle_b, fparams, bad):
    """bad: bool[ns, H, W]. -> list of first_divergence dicts (with pixel, sample)."""
    out = []
    W = fparams.width
    for s, y, x in zip(*np.nonzero(bad)):
        d = first_divergence(oracle_a, oracle_b, fparams, int(y) * W + int(x), int(s) + fparams.sample_begin)
        d.update(pixel=(int(x), int(y)), sample=int(s) + fparams.sample_begin)
        out.append(d)
    return out
