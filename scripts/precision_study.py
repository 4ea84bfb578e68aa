"""CPU emulation of the split-fp16 attention with subsets of the hi/lo product terms: how far
does the 12-layer confidence drift from the fp32 oracle?  (design study, not a test)
python scripts/precision_study.py [N] [family]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import restate, synth

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1500
family = sys.argv[2] if len(sys.argv) > 2 else "3dmatch"
cfg = synth.CONFIGS[family]
L = 12
sd = synth.make_state_dict(1234, num_layers=L, sigma_d=cfg["sigma_d"])
pair = synth.make_pair(7, N, **cfg)
_, compat = restate.spatial_compatibility(pair["src_keypts"], pair["tgt_keypts"], sd["sigma_spat"])


def split(x):
    hi = x.half().float()
    return hi, (x - hi).half().float()


def make(qk_terms, pv_terms):
    def msg(q, k, v, compat):
        C = q.shape[1]
        qs = q[0].T * (1.4426950408889634 / C ** 0.5)      # [N,C], log2 units
        qh, ql = split(qs)
        kh, kl = split(k[0].T)
        S = 0
        if "hh" in qk_terms: S = S + qh @ kh.T
        if "hl" in qk_terms: S = S + qh @ kl.T
        if "lh" in qk_terms: S = S + ql @ kh.T
        x = compat[0] * S
        m = x.max(dim=1, keepdim=True)[0]
        p = torch.exp2(x - m)
        ph, pl = split(p)
        vh, vl = split(v[0].T)                              # [N,C]
        O = 0
        if "hh" in pv_terms: O = O + ph @ vh
        if "hl" in pv_terms: O = O + ph @ vl
        if "lh" in pv_terms: O = O + pl @ vh
        l = (ph.sum(1, keepdim=True) if "rowsum_hi" in pv_terms else p.sum(1, keepdim=True))
        return (O / l).T[None]
    return msg


orig = restate.nonlocal_message
feat = restate.encoder(sd, pair["corr_pos"], compat, L)
conf0 = restate.classification(sd, feat)
print(f"N={N} {family}: conf range [{conf0.min():.3f}, {conf0.max():.3f}]")
full = ("hh", "hl", "lh")
cases = [("QK 3 / PV 3", full, full), ("QK 3 / PV hh+hl (no P_lo)", full, ("hh", "hl")),
         ("QK 3 / PV hh+hl, rowsum of P_hi", full, ("hh", "hl", "rowsum_hi")),
         ("QK 3 / PV hh+lh (no V_lo)", full, ("hh", "lh")), ("QK 3 / PV hh", full, ("hh",)),
         ("QK 3 / PV hh, rowsum of P_hi", full, ("hh", "rowsum_hi")),
         ("QK hh+hl / PV 3", ("hh", "hl"), full), ("QK hh+lh / PV 3", ("hh", "lh"), full), ("QK hh / PV 3", ("hh",), full),
         ("QK hh / PV hh", ("hh",), ("hh",))]
for name, a, b in cases:
    restate.nonlocal_message = make(a, b)
    f = restate.encoder(sd, pair["corr_pos"], compat, L)
    c = restate.classification(sd, f)
    print(f"{name:36s} max|dconf| {float((c - conf0).abs().max()):.3e}   max|dfeat| {float((f - feat).abs().max()):.3e}", flush=True)
restate.nonlocal_message = orig
