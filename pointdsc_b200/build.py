"""Builds libpointdsc_b200.so (sm_100a) in-tree with nvcc.  `python -m pointdsc_b200.build`.

Two libraries come out of the same sources:
* ``lib/libpointdsc_b200.so``         the product: no environment variable is read, every launch shape is the default;
* ``lib/libpointdsc_b200_tuning.so``  built with ``-DPDSC_TUNING``: the debug / tuning knobs (``PDSC_ATTN_GRID``,
  ``PDSC_ATTN_CLUSTER``, ``PDSC_ATTN_CHAIN``, ``PDSC_ATTN_DBG``, ``PDSC_ATTN_QK_TERMS``, ``PDSC_ATTN_PV_TERMS``,
  ``PDSC_ATTN_TRACE``, ``PDSC_KNN_FP32``) are read per call.  Only tests and scripts/ load it
  (``_cabi.use_tuning_library``).  Sources that do not mention a knob share one object file between the two."""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libpointdsc_b200.so")
TUNING_LIB_PATH = os.path.join(LIB_DIR, "libpointdsc_b200_tuning.so")
OBJ_DIR = os.path.join(HERE, "build")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
CFLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fno-gnu-unique", "--expt-relaxed-constexpr",
          "-Xptxas", "-v"]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _has_knobs(src):
    with open(src) as f:
        text = f.read()
    return any(w in text for w in ("tuning_int", "tuning_str", "PDSC_TUNING", "kTuningBuild"))


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(LIB_DIR, exist_ok=True)
    os.makedirs(OBJ_DIR, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(os.path.dirname(HERE), "include", "pointdsc_b200.h"))
    objs, tuning_objs, jobs = [], [], []
    for src in sources():
        stem = os.path.basename(src)[:-3]
        obj = os.path.join(OBJ_DIR, stem + ".o")
        objs.append(obj)
        if force or _stale(obj, [src] + headers):
            jobs.append((src, obj, []))
        if _has_knobs(src):
            tobj = os.path.join(OBJ_DIR, stem + ".tuning.o")
            tuning_objs.append(tobj)
            if force or _stale(tobj, [src] + headers):
                jobs.append((src, tobj, ["-DPDSC_TUNING"]))
        else:
            tuning_objs.append(obj)

    def compile_one(job):
        src, obj, defs = job
        with open(src) as f:
            first = f.readline()
        extra = first.split("nvcc-flags:", 1)[1].split() if "nvcc-flags:" in first else []   # per-file flags (first line)
        cmd = [NVCC] + ARCH_FLAGS + CFLAGS + extra + defs + ["-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
        log = r.stdout + r.stderr
        with open(obj[:-2] + ".ptxas.log", "w") as f:
            f.write(log)
        if verbose:
            print(log)
        return obj

    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(compile_one, jobs))
    for path, members in ((LIB_PATH, objs), (TUNING_LIB_PATH, tuning_objs)):
        if jobs or force or _stale(path, members):
            cmd = [NVCC] + ARCH_FLAGS + ["-shared", "-o", path] + members + ["-lcudart"]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
