"""Times pdsc_mlp_chain variants alone (rows = 16 x 5000): python scripts/prof_mlp.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pointdsc_b200 import _cabi
rows = 80000
g = torch.Generator().manual_seed(0)
x = torch.randn(rows, 128, generator=g).cuda()
r = torch.randn(rows, 128, generator=g).cuda()
def mk(dims, relu):
    return _cabi.MlpChain([((torch.randn(dims[i], dims[i + 1], generator=g) * 0.1).cuda(), torch.randn(dims[i + 1], generator=g).cuda()) for i in range(len(relu))], relu)
front, back = mk((128, 128), (1,)), mk((128, 64, 64, 128), (1, 1, 0))
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def t(name, fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    print(f"{name:48s} {1e3 * tot / reps:8.1f} us")
t("front y+packed", lambda: _cabi.mlp_chain(x, front, want_y=True, want_packed=True))
t("front y only", lambda: _cabi.mlp_chain(x, front, want_y=True, want_packed=False))
t("front packed only", lambda: _cabi.mlp_chain(x, front, want_y=False, want_packed=True))
t("back y (+add_out)", lambda: _cabi.mlp_chain(x, back, add_out=r))
t("back y (no add_out)", lambda: _cabi.mlp_chain(x, back))
t("torch copy 41MB (x.clone)", lambda: x.clone())
