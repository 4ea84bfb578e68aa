"""Runs the attention kernel alone on the bench shape (for ncu): python scripts/prof_attn.py [mode] [B] [N] [reps]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pointdsc_b200 import _cabi
mode = {"fp32": 0, "split16": 1, "fp16": 2, "split16pv2": 3}[sys.argv[1] if len(sys.argv) > 1 else "split16"]
if any(k.startswith("PDSC_") for k in os.environ):      # knobs only exist in the -DPDSC_TUNING build
    _cabi.use_tuning_library().__enter__()
    print("tuning library:", {k: v for k, v in os.environ.items() if k.startswith("PDSC_")})
B = int(sys.argv[2]) if len(sys.argv) > 2 else 16
N = int(sys.argv[3]) if len(sys.argv) > 3 else 5000
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 5
g = torch.Generator().manual_seed(0)
qkv = torch.randn(B, N, 384, generator=g).cuda()
src = (torch.rand(B, N, 3, generator=g) * 3).cuda()
tgt = (torch.rand(B, N, 3, generator=g) * 3).cuda()
C = 128
import time
t_end = time.perf_counter() + (0.0 if os.environ.get("PROF_NO_WARMUP") else 1.0)      # clocks ramp up from idle over ~100 ms
n_warm = 0
while n_warm < 2 or time.perf_counter() < t_end:
    out = _cabi.nonlocal_attn(qkv[:, :, :C], qkv[:, :, C:2 * C], qkv[:, :, 2 * C:], src, tgt, 0.1, mode=mode)
    torch.cuda.synchronize(); n_warm += 1
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    out = _cabi.nonlocal_attn(qkv[:, :, :C], qkv[:, :, C:2 * C], qkv[:, :, 2 * C:], src, tgt, 0.1, mode=mode)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print(f"mode {mode} B={B} N={N}: {ms:.3f} ms per call (prep+main), {4.0 * N * N * C * B / ms / 1e9:.1f} TFLOP/s algorithmic")
if mode != 0:
    # main kernel alone, on the operand tiles the last call left in the workspace: five groups of `reps` launches
    ws = _cabi._ws(_cabi.workspace_bytes(_cabi.OP_NONLOCAL_ATTN, B, N, C, mode=mode), qkv.device, None)
    groups = []
    for _ in range(5):
        e0.record()
        for _ in range(reps):
            out2 = _cabi.nonlocal_attn_packed(ws, B, N, C, mode=mode)
        e1.record()
        torch.cuda.synchronize()
        groups.append(e0.elapsed_time(e1) / reps)
    ms = sorted(groups)[2]
    print(f"mode {mode} B={B} N={N}: {ms:.3f} ms per call (main kernel only; median of 5 groups, min {min(groups):.3f} max {max(groups):.3f}), "
          f"{4.0 * N * N * C * B / ms / 1e9:.1f} TFLOP/s algorithmic; same result as prep+main: {bool(torch.equal(out, out2))}")
