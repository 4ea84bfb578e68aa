"""CUDA-graph replay of the forward for callers that run many batches of one shape (small N especially: at 64 pairs x
N=1000 the ~54 launches of a forward leave the GPU idle between kernels for ~6 % of the time).

    fwd = GraphedForward(model, B, N)                    # three eager warm-up forwards, then one capture
    res = fwd(corr_pos, src_keypts, tgt_keypts)          # copies the inputs into static buffers, replays, returns the
                                                         # STATIC result tensors (valid until the next call)

Every entry point of the library is allocation-free given its workspace and the forward never synchronises with the host
(models/PointDSC.py:128-197 re-designed that way), so the whole forward captures as it is; results are bit-identical to the
eager call.  The reference has no counterpart (its test path is batch size 1 with host round trips)."""
from __future__ import annotations

import torch

_KEYS = ("corr_pos", "src_keypts", "tgt_keypts")


class GraphedForward:
    def __init__(self, model, B: int, N: int, device=None, in_dim: int = 6, warmup: int = 3):
        device = torch.device(device if device is not None else next(model.parameters()).device)
        if device.type != "cuda":
            raise RuntimeError("GraphedForward needs a CUDA device (pointdsc_b200 has no CPU path)")
        self.model, self.B, self.N = model, B, N
        self.static_in = {"corr_pos": torch.zeros(B, N, in_dim, device=device), "src_keypts": torch.zeros(B, N, 3, device=device),
                          "tgt_keypts": torch.zeros(B, N, 3, device=device)}
        self.graph, self.out = None, None
        self._warmup = warmup

    def _capture(self):
        data = dict(self.static_in)
        data["testing"] = True
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):                      # workspaces, packed weights and shared-memory opt-ins happen here
            for _ in range(self._warmup):
                self.model(data)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.out = self.model(data)

    def __call__(self, corr_pos, src_keypts, tgt_keypts):
        for k, t in zip(_KEYS, (corr_pos, src_keypts, tgt_keypts)):
            if t.shape != self.static_in[k].shape:
                raise ValueError(f"{k}: shape {tuple(t.shape)} does not match the captured {tuple(self.static_in[k].shape)}")
            self.static_in[k].copy_(t, non_blocking=True)
        if self.graph is None:
            self._capture()                                # (the inputs just copied take part in the warm-up forwards)
        self.graph.replay()
        return self.out
