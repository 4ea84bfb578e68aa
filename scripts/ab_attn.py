"""Times the attention main kernel of several library variants (scripts/ab_variant.py) against each other, interleaved, in
one process:   python scripts/ab_attn.py B N reps libA.so libB.so ...     (paths relative to pointdsc_b200/lib/ab/)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from pointdsc_b200 import _cabi
B, N, reps = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
libs = [os.path.join(ROOT, "pointdsc_b200", "lib", "ab", x) if not os.path.isabs(x) else x for x in sys.argv[4:]]
C, mode = 128, 1
g = torch.Generator().manual_seed(0)
qkv = torch.randn(B, N, 384, generator=g).cuda()
src = (torch.rand(B, N, 3, generator=g) * 3).cuda()
tgt = (torch.rand(B, N, 3, generator=g) * 3).cuda()
outs, times = {}, {l: [] for l in libs}
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
import time
for rnd in range(8):
    for l in libs:
        _cabi._active = l
        if rnd == 0:
            t_end = time.perf_counter() + 0.7
            while time.perf_counter() < t_end:
                outs[l] = _cabi.nonlocal_attn(qkv[:, :, :C], qkv[:, :, C:2 * C], qkv[:, :, 2 * C:], src, tgt, 0.1, mode=mode)
                torch.cuda.synchronize()
        ws = _cabi._ws(_cabi.workspace_bytes(_cabi.OP_NONLOCAL_ATTN, B, N, C, mode=mode), qkv.device, None)
        _cabi.nonlocal_attn(qkv[:, :, :C], qkv[:, :, C:2 * C], qkv[:, :, 2 * C:], src, tgt, 0.1, mode=mode)
        e0.record()
        for _ in range(reps):
            o = _cabi.nonlocal_attn_packed(ws, B, N, C, mode=mode)
        e1.record()
        torch.cuda.synchronize()
        if rnd:
            times[l].append(e0.elapsed_time(e1) / reps)
        assert torch.equal(o, outs[l])
base = outs[libs[0]]
for l in libs:
    t = sorted(times[l])
    print(f"B={B} N={N} {os.path.basename(l):28s} median {t[len(t) // 2]:.4f} ms  min {t[0]:.4f}  max {t[-1]:.4f}   "
          f"identical to {os.path.basename(libs[0])}: {bool(torch.equal(outs[l], base))}  max|diff| {float((outs[l] - base).abs().max()):.2e}")
