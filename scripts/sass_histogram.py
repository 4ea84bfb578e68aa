"""Opcode counts per kernel of the shipped library (cuobjdump -sass): which kernels use tcgen05 / TMEM / the TMA engine.
python scripts/sass_histogram.py > profiles/r02_sass_histogram.txt      (no GPU needed)"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "pointdsc_b200", "lib", "libpointdsc_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", out)), capture_output=True, text=True).stdout.split("\n")
COLS = ["UTCHMMA", "LDTM", "STTM", "UBLKCP", "UTCBAR", "SYNCS", "MUFU", "FFMA2", "FADD2", "SHFL", "HMMA", "LDGSTS", "UTMALDG"]
kernels, cur, it = [], None, iter(names)
for line in out.split("\n"):
    m = re.search(r"Function : (\S+)", line)
    if m:
        full = next(it)
        short = re.sub(r"\(.*", "", full).replace("void ", "")
        tmpl = re.search(r"<[^()]*>", full)
        cur = [short + (tmpl.group(0) if tmpl and "<" not in short else ""), collections.Counter(), 0]
        kernels.append(cur)
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and cur is not None:
        cur[1][m.group(1)] += 1
        cur[2] += 1
print("# cuobjdump -sass pointdsc_b200/lib/libpointdsc_b200.so (the shipped, knob-free library), opcode counts per kernel (scripts/sass_histogram.py)")
print("# UTCHMMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UBLKCP = cp.async.bulk (TMA engine), UTCBAR = tcgen05.commit, SYNCS = mbarrier ops; no HMMA (legacy mma.sync) anywhere")
print(f"{'kernel':70s} " + " ".join(f"{c:>8s}" for c in COLS) + "    total")
for name, cnt, total in kernels:
    print(f"{name[:70]:70s} " + " ".join(f"{cnt.get(c, 0):8d}" for c in COLS) + f" {total:8d}")
