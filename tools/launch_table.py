"""Markdown table of an ncu launch list (gpu__time_duration.sum csv):  python tools/launch_table.py launches.csv"""
import collections, csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    k = r[4].split('(')[0].replace('void ', '')
    agg[k][0] += 1; agg[k][1] += float(r[14]) / 1e3
tot = sum(v[1] for v in agg.values())
print("| kernel | launches | total us | share | avg us |\n|---|---|---|---|---|")
for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"| {k} | {v[0]} | {v[1]:.1f} | {100*v[1]/tot:.1f}% | {v[1]/v[0]:.1f} |")
