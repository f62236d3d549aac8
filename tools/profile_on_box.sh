#!/bin/bash
# Runs on the GPU box (under gpurun): ncu captures of one frame of a workload, summarised THERE so that only small files come
# back in gpurun_out/ (the full report of ~80 launches is > 100 MB; gpurun copies back at most 64 MiB).
#   tools/profile_on_box.sh <workload> <spp> <tag>        e.g.  tools/profile_on_box.sh c2 2 r02d
set -u
W=$1; SPP=$2; TAG=$3
OUT=gpurun_out; TMP=/tmp/hjprof_$W; mkdir -p $OUT $TMP
python tools/prof_frame.py $W $SPP pipelines=1 --stats-json $OUT/${TAG}_stats_$W.json > $OUT/${TAG}_prof_$W.log 2>&1 || { echo "prof_frame failed"; tail -5 $OUT/${TAG}_prof_$W.log; exit 1; }
ALGO=$(python -c "import json; print(json.load(open('$OUT/${TAG}_stats_$W.json'))['algorithmic_bytes_per_launch_80B'])")
python tools/prof_frame.py $W $SPP pipelines=1 --one-frame > $OUT/${TAG}_plain_$W.log 2>&1 || { echo "plain run failed"; exit 1; }
# every traversal + shading launch of the frame, full metric set, no source (kept on the box)
ncu --set full --clock-control none -k regex:"k_trace|k_surface|k_nee|k_sample" -o $TMP/frame python tools/prof_frame.py $W $SPP pipelines=1 --one-frame > $OUT/${TAG}_ncu_$W.log 2>&1
python tools/ncu_summary.py $TMP/frame.ncu-rep --json $OUT/${TAG}_ncu_trace_$W.json --algo-bytes-per-launch $ALGO \
    --note "ncu --set full --clock-control none over every traversal and shading launch of one frame: tools/prof_frame.py $W $SPP pipelines=1 --one-frame" > /dev/null
ncu -i $TMP/frame.ncu-rep --page raw --csv > $TMP/raw.csv 2>/dev/null
python - <<PY
import csv
rows = list(csv.reader(open("$TMP/raw.csv")))
keep = ['ID', 'Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct', 'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__issue_active.avg.per_cycle_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'smsp__inst_executed.sum', 'launch__grid_size',
        'sm__inst_executed_pipe_alu.sum', 'sm__inst_executed_pipe_fma.sum', 'sm__inst_executed_pipe_fp64.sum', 'sm__inst_executed_pipe_lsu.sum', 'sm__inst_executed_pipe_xu.sum',
        'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active']
hdr = rows[0]
idx = [i for i, h in enumerate(hdr) if h in keep or ('issue_stalled' in h and h.endswith('_per_issue_active.ratio'))]
with open("$OUT/${TAG}_ncu_launches_$W.csv", "w", newline="") as f:
    w = csv.writer(f)
    for r in rows:
        w.writerow([r[i][:70] for i in idx])
PY
# the first traversal launches with source (small report, comes back)
ncu --set full --clock-control none --import-source on -k regex:"k_trace" -c 3 -o $OUT/${TAG}_k_trace_$W python tools/prof_frame.py $W $SPP pipelines=1 --one-frame >> $OUT/${TAG}_ncu_$W.log 2>&1
if [ "$W" = "c2" ]; then
  ncu --set full --clock-control none --import-source on -k regex:"k_surface|k_nee|k_sample" -c 5 -o $OUT/${TAG}_shading_$W python tools/prof_frame.py $W $SPP pipelines=1 --one-frame >> $OUT/${TAG}_ncu_$W.log 2>&1
fi
ls -la $OUT | tail -12
