"""Summarises an `ncu --set full` report of a frame (tools/prof_frame.py) per kernel group.

    python tools/ncu_summary.py report.ncu-rep                    # the handful of metrics the profile notes quote, per launch
    python tools/ncu_summary.py report.ncu-rep --json out.json [--algo-bytes-per-launch B] [--note "..."]
        per kernel group (traversal = k_trace*, shading = k_surface / k_nee / k_sample): launches, time, DRAM bytes per launch,
        DRAM GB/s, L1 / L2 hit rates, issue slots per active cycle, lanes per instruction, top stalls — duration-weighted.
        bench.py reads profiles/r02_ncu_trace_<workload>.json written this way for roofline.traffic."""
import csv
import io
import json
import subprocess
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "ns": 1e-9, "us": 1e-6, "usecond": 1e-6, "ms": 1e-3, "msecond": 1e-3,
        "s": 1.0, "second": 1.0, "nsecond": 1e-9}


def load(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    return rows[0], rows[1], rows[2:]


def main():
    rep = sys.argv[1]
    hdr, units, rows = load(rep)
    col = {h: i for i, h in enumerate(hdr)}

    def val(r, name, default=0.0):
        if name not in col:
            return default
        try:
            return float(r[col[name]].replace(",", "")) * UNIT.get(units[col[name]], 1.0)
        except ValueError:
            return default

    if "--json" not in sys.argv:
        want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
                'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
                'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__issue_active.avg.per_cycle_active',
                'sm__cycles_active.avg', 'sm__cycles_elapsed.avg', 'sm__warps_active.avg.pct_of_peak_sustained_active',
                'launch__registers_per_thread', 'smsp__inst_executed.sum', 'smsp__average_warp_latency_per_inst_issued.ratio',
                'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'launch__grid_size']
        for w in want:
            if w in col:
                print(w, units[col[w]], [r[col[w]][:60] for r in rows])
        for h, i in col.items():
            if 'issue_stalled' in h and h.endswith('_per_issue_active.ratio'):
                vals = [r[i] for r in rows]
                try:
                    if max(float(v) for v in vals) > 0.4:
                        print(h.replace('smsp__average_warps_issue_stalled_', 'stall ').replace('_per_issue_active.ratio', ''), vals)
                except ValueError:
                    pass
        return

    out_path = sys.argv[sys.argv.index("--json") + 1]
    algo = float(sys.argv[sys.argv.index("--algo-bytes-per-launch") + 1]) if "--algo-bytes-per-launch" in sys.argv else None
    note = sys.argv[sys.argv.index("--note") + 1] if "--note" in sys.argv else None
    groups = {"traversal": lambda n: "k_trace" in n, "shading": lambda n: any(k in n for k in ("k_surface", "k_nee", "k_sample")),
              "k_surface": lambda n: "k_surface" in n, "k_nee": lambda n: "k_nee" in n, "k_sample": lambda n: "k_sample" in n}
    stalls = [h for h in hdr if 'issue_stalled' in h and h.endswith('_per_issue_active.ratio')]
    out = {"report": rep}
    for g, pred in groups.items():
        rs = [r for r in rows if pred(r[col['Kernel Name']])]
        if not rs:
            continue
        t = [val(r, 'gpu__time_duration.sum') for r in rs]
        T = sum(t) or 1.0
        wavg = lambda name: sum(val(r, name) * ti for r, ti in zip(rs, t)) / T
        dram = sum(val(r, 'dram__bytes_read.sum') + val(r, 'dram__bytes_write.sum') for r in rs)
        d = dict(launches=len(rs), time_ms=round(1e3 * T, 3), avg_launch_us=round(1e6 * T / len(rs), 1),
                 dram_bytes_per_launch=round(dram / len(rs)), dram_gbs=round(dram / T / 1e9, 1),
                 l1_hit=round(wavg('l1tex__t_sector_hit_rate.pct') / 100, 4), l2_hit=round(wavg('lts__t_sector_hit_rate.pct') / 100, 4),
                 issue_active=round(wavg('smsp__issue_active.avg.per_cycle_active'), 3),
                 lanes_per_inst=round(wavg('smsp__thread_inst_executed_per_inst_executed.ratio'), 2),
                 warps_active_pct=round(wavg('sm__warps_active.avg.pct_of_peak_sustained_active'), 1),
                 registers=int(val(rs[0], 'launch__registers_per_thread')),
                 dram_pct_of_peak=round(wavg('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'), 1),
                 stalls_per_issue={h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''): round(wavg(h), 2)
                                   for h in stalls if wavg(h) > 0.4})
        if g == "traversal":
            if algo:
                d["algorithmic_bytes_per_launch"] = round(algo)
                d["traffic_over_algorithmic"] = round(dram / len(rs) / algo, 4)
            if note:
                d["note"] = note
            d["bound_by"] = ("issue / latency (%.2f issue slots per active cycle and scheduler, %.1f of 32 lanes per instruction, DRAM at %.0f %% of peak)"
                             % (d["issue_active"], d["lanes_per_inst"], d["dram_pct_of_peak"]))
        out[g] = d
    json.dump(out, open(out_path, "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
