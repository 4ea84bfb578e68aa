"""Which key tile is mishandled?  For the rows that are wrong, compare with reference variants (tile dropped / duplicated)."""
import os, sys, itertools
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from oracle import restate
from pointdsc_b200 import _cabi
N = int(sys.argv[1]); grid = sys.argv[2]; B = 2
g = torch.Generator().manual_seed(1000 + N)
C = 128
qkv = torch.randn(B, N, 3 * C, generator=g); src = torch.rand(B, N, 3, generator=g); tgt = src + 0.05 * torch.randn(B, N, 3, generator=g)
d = qkv.cuda()
os.environ["PDSC_ATTN_GRID"] = grid
with _cabi.use_tuning_library():
    out = _cabi.nonlocal_attn(d[:, :, :C], d[:, :, C:2*C], d[:, :, 2*C:], src.cuda(), tgt.cuda(), 0.1, mode=_cabi.ATTN_SPLIT16).cpu()
NT = (N + 63) // 64
for b in range(B):
    _, compat = restate.spatial_compatibility(src[b:b+1], tgt[b:b+1], torch.tensor([0.1]))
    q, k, v = qkv[b, :, :C].double(), qkv[b, :, C:2*C].double(), qkv[b, :, 2*C:].double()
    x = compat[0].double() * (q @ k.T) / C ** 0.5
    mx = x.max(dim=1, keepdim=True)[0]
    p = torch.exp(x - mx)
    Ot = torch.stack([p[:, t*64:(t+1)*64] @ v[t*64:(t+1)*64] for t in range(NT)])     # [NT,N,C]
    lt = torch.stack([p[:, t*64:(t+1)*64].sum(1) for t in range(NT)])               # [NT,N]
    ref = Ot.sum(0) / lt.sum(0)[:, None]
    err = (out[b].double() - ref).abs().amax(-1)
    for r0 in range(0, N, 128):
        rows = slice(r0, min(N, r0 + 128))
        if float(err[rows].max()) < 1e-4: continue
        bad = int((err[rows] > 1e-4).sum())
        best = []
        for w in itertools.product([0, 1, 2], repeat=NT) if NT <= 8 else []:
            if sum(1 for a in w if a != 1) > 2: continue
            wt = torch.tensor(w, dtype=torch.float64)
            var = (wt[:, None, None] * Ot[:, rows]).sum(0) / (wt[:, None] * lt[:, rows]).sum(0)[:, None]
            best.append((float((out[b, rows].double() - var).abs().max()), w))
        best.sort()
        print(f"pair {b} rows {r0}..: {bad} bad rows, max err {float(err[rows].max()):.2e}; best tile-weight explanations: {best[:3]}", flush=True)
