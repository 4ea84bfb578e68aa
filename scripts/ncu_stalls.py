"""Source-page stall summary of an ncu report (--set full --import-source on): totals per stall reason and the instructions with
the most samples.   python scripts/ncu_stalls.py report.ncu-rep [top]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
kernel, hdr, body = rows[0][1], rows[1], rows[2:]
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
ns = hdr.index("# Samples")
base = int(body[0][0], 16)
tot = {hdr[i]: 0 for i in stall}
recs = []
for r in body:
    if len(r) <= max(stall):
        continue
    per = {hdr[i]: int(r[i] or 0) for i in stall}
    for k, v in per.items():
        tot[k] += v
    recs.append((int(r[ns] or 0), int(r[0], 16) - base, r[1].strip(), per))
total = sum(tot.values())
print("kernel:", kernel)
print("total warp-stall samples:", total)
for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:10]:
    print(f"  {k:24s} {v:8d}  {100.0 * v / total:4.1f}%")
print("top instructions by samples (address, samples, SASS, top two stall reasons):")
for n, addr, sass, per in sorted(recs, key=lambda x: -x[0])[:top]:
    two = sorted(per.items(), key=lambda kv: -kv[1])[:2]
    print(f"  {addr:05x} {n:7d}  {sass[:70]:70s} " + " ".join(f"{k[6:]}:{v}" for k, v in two))
