"""Groups the SASS page of an ncu report into runs of equal execution count (≈ basic blocks) and prints each run's share
of the warp instructions and its average active lanes:  python tools/ncu_sass_blocks.py rep.ncu-rep kernel_regex [launch]"""
import csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern, "--launch-skip", skip,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
his = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
hi = his[0]
hdr = rows[hi]
end = his[1] - 1 if len(his) > 1 else len(rows)
data = [r for r in rows[hi + 1:end] if len(r) == len(hdr)]
ia, it, isrc = hdr.index('Instructions Executed'), hdr.index('Thread Instructions Executed'), hdr.index('Source')
ist = hdr.index('# Samples')
tot = sum(int(r[ia]) for r in data); tt = sum(int(r[it]) for r in data); ts = sum(int(r[ist]) for r in data)
print(rows[hi - 1][1] if len(rows[hi - 1]) > 1 else "", "\ntotal warp inst", tot, "thread inst", tt, "lanes %.2f" % (tt / tot), "samples", ts)
runs, run, prev = [], [], None
for k, r in enumerate(data):
    c = int(r[ia])
    if prev is None or abs(c - prev) > 0.02 * max(c, prev, 1):
        if run: runs.append(run)
        run = []
    run.append((k, r[isrc].split()[0] if r[isrc] else '', c, int(r[it]), int(r[ist]), r[isrc])); prev = c
runs.append(run)
for run in runs:
    c = sum(x[2] for x in run); t = sum(x[3] for x in run); s = sum(x[4] for x in run)
    if c / tot > 0.004:
        ops = {}
        for x in run:
            o = x[1].split('.')[0]
            if o.startswith('@'): o = x[5].split()[1].split('.')[0] if len(x[5].split()) > 1 else o
            ops[o] = ops.get(o, 0) + 1
        top = sorted(ops.items(), key=lambda z: -z[1])[:7]
        print(f"[{run[0][0]:4d}-{run[-1][0]:4d}] n={len(run):4d} inst={100*c/tot:5.1f}% samples={100*s/max(ts,1):5.1f}% lanes={t/max(c,1):5.1f} exec={run[0][2]:9d} {top}")
