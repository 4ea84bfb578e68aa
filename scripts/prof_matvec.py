"""compat_matvec on the row range a rank owns at world sizes 1/2/4/8 (one GPU is enough: the kernel does not know about
peers), for every column split: python scripts/prof_matvec.py [N]   (tuning library: PDSC_MV_KS)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import synth
from pointdsc_b200 import _cabi
N = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
cfg = synth.CONFIGS["kitti"]
p = synth.make_pair(200, N, **cfg)
s, t = p["src_keypts"][0].cuda().contiguous(), p["tgt_keypts"][0].cuda().contiguous()
v = torch.rand(N, device="cuda")
def timeit(fn, reps=20):
    t_end = time.perf_counter() + 0.3
    while time.perf_counter() < t_end: fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3
with _cabi.use_tuning_library():
    flag = torch.zeros(1, dtype=torch.int32, device="cuda"); it = torch.zeros(1, dtype=torch.int32, device="cuda"); y = torch.rand(N, device="cuda")
    print(f"N={N}: power_normalize {timeit(lambda: (flag.zero_(), _cabi.power_normalize(y, v.clone(), flag, it))):.1f} us (incl. a clone and a memset)")
    for ks in ("4", "5", "6", "7", "8"):
        os.environ["PDSC_MV_KS"] = ks
        row = []
        for P in (1, 2, 4, 8):
            rows = (N + P - 1) // P
            out = torch.empty(rows, device="cuda")
            us = timeit(lambda: _cabi.compat_matvec(s, t, v, 0, rows, cfg["sigma_d"], out=out))
            row.append(f"1/{P}: {us:7.1f} us (x{P} = {us * P:7.1f})")
        print(f"column split {ks}: " + "   ".join(row), flush=True)
