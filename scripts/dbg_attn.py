"""Debug aid: attention vs the oracle for given N and grid settings, error per 128-row tile, run twice (determinism)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from oracle import restate
from pointdsc_b200 import _cabi
N = int(sys.argv[1]); grids = sys.argv[2].split(","); B = int(sys.argv[3]) if len(sys.argv) > 3 else 2
g = torch.Generator().manual_seed(1000 + N)
C = 128
qkv = torch.randn(B, N, 3 * C, generator=g); src = torch.rand(B, N, 3, generator=g); tgt = src + 0.05 * torch.randn(B, N, 3, generator=g)
refs = []
for b in range(B):
    _, compat = restate.spatial_compatibility(src[b:b+1], tgt[b:b+1], torch.tensor([0.1]))
    refs.append(restate.nonlocal_message(qkv[b:b+1, :, :C].permute(0,2,1), qkv[b:b+1, :, C:2*C].permute(0,2,1), qkv[b:b+1, :, 2*C:].permute(0,2,1), compat).permute(0,2,1))
ref = torch.cat(refs)
d = qkv.cuda()
with _cabi.use_tuning_library():
    for grid in grids:
        if grid == "none": os.environ.pop("PDSC_ATTN_GRID", None)
        else: os.environ["PDSC_ATTN_GRID"] = grid
        outs = []
        for rep in range(3):
            out = _cabi.nonlocal_attn(d[:, :, :C], d[:, :, C:2*C], d[:, :, 2*C:], src.cuda(), tgt.cuda(), 0.1, mode=_cabi.ATTN_SPLIT16)
            torch.cuda.synchronize(); outs.append(out.cpu())
        err = (outs[0] - ref).abs().amax(dim=-1)      # [B,N]
        tiles = [[float(err[b, i:i+128].max()) for i in range(0, N, 128)] for b in range(B)]
        print(f"N={N} grid={grid}: max err {float(err.max()):.2e}; deterministic {all(torch.equal(outs[0], o) for o in outs[1:])}; per 128-row tile: " +
              " | ".join(" ".join(f"{e:.0e}" for e in tb) for tb in tiles), flush=True)
