"""Summarise an ncu launch list (gpu__time_duration.sum per launch): python scripts/summarize_launches.py launches.csv [last_n]"""
import csv, collections, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi = hdr.index('Kernel Name'), hdr.index('Metric Value')
seq = []
for r in rows[1:]:
    try:
        seq.append((r[ki], float(r[vi].replace(',', ''))))
    except ValueError:
        pass
last = int(sys.argv[2]) if len(sys.argv) > 2 else len(seq)
seq = seq[-last:]
agg = collections.defaultdict(lambda: [0, 0.0])
for k, v in seq:
    agg[k[:100]][0] += 1
    agg[k[:100]][1] += v
tot = sum(v for _, v in seq)
print(f"# {len(seq)} launches, total {tot/1e3:.1f} us")
print("kernel,launches,total_ns,share")
for k, (c, v) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"\"{k}\",{c},{v:.0f},{v/tot:.4f}")
