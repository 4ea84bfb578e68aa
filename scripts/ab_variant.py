"""A/B helper: builds a variant of ONE kernel file into pointdsc_b200/lib/ab/NAME.so (the other objects are the product
build's), so that scripts/ab_attn.py can time variants against each other inside one process on one box -- timings from
different boxes of the pool differ by +-5 %, more than most of the changes worth measuring.
    python scripts/ab_variant.py NAME [--ref GITREF] [--file attn_pp.cu] [-DFLAG ...]
--ref takes the file from a git revision instead of the working copy."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pointdsc_b200 import build as B

name = sys.argv[1]
ref, fname, defs = None, "attn_pp.cu", []
a = sys.argv[2:]
while a:
    x = a.pop(0)
    if x == "--ref": ref = a.pop(0)
    elif x == "--file": fname = a.pop(0)
    else: defs.append(x)
B.build()
src = os.path.join(B.CSRC, fname)
if ref:
    text = subprocess.run(["git", "show", f"{ref}:pointdsc_b200/csrc/{fname}"], cwd=ROOT, capture_output=True, text=True, check=True).stdout
    src = os.path.join(B.OBJ_DIR, f"ab_{name}_{fname}")
    with open(src, "w") as f:
        f.write(text)
out_dir = os.path.join(B.LIB_DIR, "ab")
os.makedirs(out_dir, exist_ok=True)
obj = os.path.join(B.OBJ_DIR, f"ab_{name}.o")
cmd = [B.NVCC] + B.ARCH_FLAGS + B.CFLAGS + ["-I", B.CSRC] + defs + ["-c", src, "-o", obj]
r = subprocess.run(cmd, capture_output=True, text=True)
if r.returncode:
    sys.exit(r.stdout + r.stderr)
stem = fname[:-3]
objs = [os.path.join(B.OBJ_DIR, os.path.basename(s)[:-3] + ".o") for s in B.sources() if os.path.basename(s)[:-3] != stem] + [obj]
lib = os.path.join(out_dir, name + ".so")
subprocess.run([B.NVCC] + B.ARCH_FLAGS + ["-shared", "-o", lib] + objs + ["-lcudart"], check=True)
print(lib)
