"""ctypes binding of libpointdsc_b200.so (include/pointdsc_b200.h).

This is the stub a maintainer of the reference would add: raw device
pointers and sizes in, an int status out; torch only supplies ``data_ptr()`` and the
current stream.  There is no CPU fallback: a missing library or a non-CUDA tensor raises.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_double, c_float, c_int, c_int64, c_size_t, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libpointdsc_b200.so")
# the same sources built with -DPDSC_TUNING: the only build that reads the PDSC_* debug / tuning environment variables
# (tests and scripts/ select it with use_tuning_library(); the product path never loads it)
TUNING_LIB_PATH = os.path.join(_HERE, "lib", "libpointdsc_b200_tuning.so")

ATTN_FP32, ATTN_SPLIT16, ATTN_FP16, ATTN_SPLIT16_PV2 = 0, 1, 2, 3
OP_NONLOCAL_ATTN, OP_PICK_SEEDS, OP_SEED_KNN, OP_SEED_HYPOTHESES = 1, 2, 3, 4
OP_SCORE_HYPOTHESES, OP_POST_REFINE, OP_POWER_ITER_FULL = 5, 6, 7
OP_HYBRID_SEARCH, OP_VOXEL_DOWN_SAMPLE, OP_NN_MATCH_F64, OP_ICP_REFINE = 8, 9, 10, 11

# name -> (restype, argtypes); must list every symbol the header declares
SIGNATURES = {
    "pdsc_version": (c_int, []),
    "pdsc_last_error": (ctypes.c_char_p, []),
    "pdsc_check_device": (c_int, []),
    "pdsc_launch_count": (ctypes.c_uint64, []),
    "pdsc_workspace_bytes": (c_size_t, [c_int] * 7),
    "pdsc_nonlocal_attn": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64,
                                   c_int, c_int, c_int, c_float, c_int, c_void_p, c_size_t, c_void_p]),
    "pdsc_pick_seeds": (c_int, [c_void_p, c_void_p, c_int, c_int, c_float, c_int, c_void_p, c_void_p, c_void_p,
                                c_void_p, c_size_t, c_void_p]),
    "pdsc_seed_knn": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                              c_void_p, c_size_t, c_void_p]),
    "pdsc_seed_hypotheses": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
                                     c_float, c_float, c_int, c_void_p, c_void_p, c_void_p,
                                     c_void_p, c_size_t, c_void_p]),
    "pdsc_score_hypotheses": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float,
                                      c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "pdsc_post_refine": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_int, c_void_p, c_void_p,
                                 c_void_p]),
    "pdsc_feature_head": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_float, c_void_p, c_void_p, c_void_p]),
    "pdsc_split16_rows": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_void_p]),
    "pdsc_qkv_pack": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_void_p, c_size_t,
                              c_void_p]),
    "pdsc_nonlocal_attn_packed": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_int, c_void_p, c_size_t, c_void_p]),
    "pdsc_nn_match": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                              c_void_p]),
    "pdsc_mutual_corr": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "pdsc_registration_metrics": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "pdsc_classification_counts": (c_int, [c_void_p, c_void_p, c_int, c_int, c_float, c_void_p, c_void_p, c_void_p]),
    "pdsc_mlp_chain": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                               c_void_p, c_void_p, c_int, c_void_p]),
    "pdsc_mlp_chain_direct": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                      c_void_p, c_void_p]),
    "pdsc_compat_matvec": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float,
                                   c_void_p, c_void_p]),
    "pdsc_power_normalize": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "pdsc_hybrid_search": (c_int, [c_void_p, c_int, c_double, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                   c_void_p]),
    "pdsc_voxel_down_sample": (c_int, [c_void_p, c_void_p, c_int, c_double, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                       c_void_p]),
    "pdsc_fpfh": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "pdsc_nn_match_f64": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "pdsc_icp_refine": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_double, c_int, c_double, c_double, c_void_p, c_void_p,
                                c_void_p, c_void_p, c_size_t, c_void_p]),
    "pdsc_set_overflow_flag": (c_int, [c_void_p]),
    "pdsc_power_p2p_exchange_bytes": (c_size_t, [c_int]),
    "pdsc_compat_matvec_p2p": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_void_p, c_void_p, c_int, c_int,
                                       c_int, c_void_p]),
    "pdsc_power_normalize_p2p": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "pdsc_power_iter_full_static": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_float, c_int, c_void_p, c_void_p]),
    "pdsc_power_iter_full": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_float, c_int, c_void_p,
                                     c_void_p, c_size_t, c_void_p]),
}

_libs = {}         # path -> loaded library
_active = LIB_PATH
launch_count = 0   # kernels-launching C-ABI calls made through this binding (bench.py reports it)


def load():
    """Loads the shared library (no CUDA context needed) and declares every signature."""
    lib = _libs.get(_active)
    if lib is None:
        if not os.path.exists(_active):
            raise RuntimeError(
                f"{_active} is missing: build it with `python -m pointdsc_b200.build` "
                "(there is no CPU or PyTorch fallback for the CUDA path)")
        lib = ctypes.CDLL(_active)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _libs[_active] = lib
    return lib


class use_tuning_library:
    """Context manager for tests / scripts: calls made inside go to libpointdsc_b200_tuning.so, which reads the PDSC_*
    knobs from the environment.  The product library ignores them."""

    def __enter__(self):
        global _active
        self._prev = _active
        _active = TUNING_LIB_PATH
        return load()

    def __exit__(self, *exc):
        global _active
        _active = self._prev
        return False


def is_tuning_library():
    return _active == TUNING_LIB_PATH


def _check(rc, what):
    if rc != 0:
        msg = load().pdsc_last_error().decode(errors="replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")


def _ptr(t, dtype=None, name="tensor"):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (pointdsc_b200 has no CPU path)")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f"{name} must be {dtype}, got {t.dtype}")
    return t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


class Workspace:
    """Caller-owned scratch: grows on demand, reused across calls on one device."""

    def __init__(self):
        self.buf = None

    def get(self, nbytes, device):
        nbytes = max(int(nbytes), 256)
        if self.buf is None or self.buf.numel() < nbytes or self.buf.device != device:
            self.buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
        return self.buf


_default_ws = {}


def _ws(nbytes, device, ws):
    if ws is None:
        ws = _default_ws.setdefault(device, Workspace())
    return ws.get(nbytes, device)


def workspace_bytes(op, B=1, N=0, C=128, S=0, k=0, mode=0):
    return int(load().pdsc_workspace_bytes(op, B, N, C, S, k, mode))


def kernel_launches():
    """CUDA kernels enqueued by the library so far (counted inside the library)."""
    return int(load().pdsc_launch_count())


def check_device():
    _check(load().pdsc_check_device(), "pdsc_check_device")


def _count():
    global launch_count
    launch_count += 1


# ------------------------------------------------------------------------------------------------
def nonlocal_attn(q, k, v, src, tgt, sigma_spat, mode=ATTN_FP32, out=None, ws=None):
    """q,k,v: [B,N,C] fp32 views with a common row stride (e.g. slices of a [B,N,3C] tensor)."""
    lib = load()
    B, N, C = q.shape
    for t in (q, k, v):
        if not (t.is_cuda and t.dtype == torch.float32 and t.stride(2) == 1 and t.stride(0) == N * t.stride(1)
                and t.stride(1) == q.stride(1) and t.shape == q.shape):
            raise RuntimeError("q,k,v must be CUDA fp32 [B,N,C] with unit channel stride and equal row stride")
    if out is None:
        out = torch.empty(B, N, C, dtype=torch.float32, device=q.device)
    elif not (out.is_cuda and out.dtype == torch.float32 and out.shape == q.shape and out.stride(2) == 1
              and out.stride(0) == N * out.stride(1)):
        raise RuntimeError("out must be CUDA fp32 [B,N,C] with unit channel stride (rows may be wider than C: a view of [B,N,ld])")
    nbytes = workspace_bytes(OP_NONLOCAL_ATTN, B, N, C, mode=mode)
    w = _ws(nbytes, q.device, ws)
    rc = lib.pdsc_nonlocal_attn(q.data_ptr(), k.data_ptr(), v.data_ptr(), q.stride(1),
                                _ptr(src, torch.float32, "src"), _ptr(tgt, torch.float32, "tgt"),
                                out.data_ptr(), out.stride(1), B, N, C, float(sigma_spat), int(mode),
                                w.data_ptr(), w.numel(), _stream())
    _check(rc, "pdsc_nonlocal_attn")
    _count()
    return out


def split_kp(K):
    """Packed row length (fp16 elements) for K input channels: [hi|hi|lo|1 1 0..] rounded up to 8."""
    return (3 * K + 2 + 7) // 8 * 8


def split16_rows(x, add=None, relu=False, y_out=None, out16=None):
    """x (+add), optional ReLU -> packed fp16 [rows, split_kp(K)]; y_out (may alias x) gets the fp32 value."""
    lib = load()
    rows, K = x.shape
    Kp = split_kp(K)
    if out16 is None:
        out16 = torch.empty(rows, Kp, dtype=torch.float16, device=x.device)
    rc = lib.pdsc_split16_rows(_ptr(x, torch.float32, "x"), _ptr(add, torch.float32, "add"), _ptr(y_out, torch.float32, "y_out"),
                               _ptr(out16, torch.float16, "out16"), rows, K, Kp, 1 if relu else 0, _stream())
    _check(rc, "pdsc_split16_rows")
    _count()
    return out16


def qkv_pack(x, wpack, bias, src, tgt, sigma_spat, ws=None, mode=ATTN_SPLIT16):
    """x [B,N,128] fp32 -> attention operand tiles in the workspace (q/k/v projections fused with the packing).
    wpack = cat(pack_umma_b(Wq^T), pack_umma_b(Wk^T), pack_umma_b(Wv^T)); bias = cat(bq, bk, bv).  Returns the
    workspace tensor to hand to nonlocal_attn_packed."""
    lib = load()
    B, N, C = x.shape
    w = _ws(workspace_bytes(OP_NONLOCAL_ATTN, B, N, C, mode=mode), x.device, ws)
    rc = lib.pdsc_qkv_pack(_ptr(x, torch.float32, "x"), _ptr(wpack, torch.uint8, "wpack"), _ptr(bias, torch.float32, "bias"),
                           _ptr(src, torch.float32, "src"), _ptr(tgt, torch.float32, "tgt"), B, N, C, float(sigma_spat),
                           w.data_ptr(), w.numel(), _stream())
    _check(rc, "pdsc_qkv_pack")
    _count()
    return w


def nonlocal_attn_packed(w, B, N, C, mode=ATTN_SPLIT16, out=None):
    """Message [B,N,C] from operand tiles prepared by qkv_pack in workspace tensor `w`."""
    lib = load()
    if out is None:
        out = torch.empty(B, N, C, dtype=torch.float32, device=w.device)
    rc = lib.pdsc_nonlocal_attn_packed(_ptr(out, torch.float32, "out"), C, B, N, C, mode, w.data_ptr(), w.numel(), _stream())
    _check(rc, "pdsc_nonlocal_attn_packed")
    _count()
    return out


class MlpChain:
    """Device-side description of a fused pointwise MLP chain (pdsc_mlp_chain): pre-packed weights + biases."""

    def __init__(self, layers, relu):
        """layers: list of (Wt [K,N] fp32 with BatchNorm folded, b [N] fp32) on the target device; relu: flags."""
        self.dims = [int(layers[0][0].shape[0])] + [int(Wt.shape[1]) for Wt, _ in layers]
        self.relu = [1 if r else 0 for r in relu]
        self.c_dims = (ctypes.c_int32 * len(self.dims))(*self.dims)
        self.c_relu = (ctypes.c_int32 * len(self.relu))(*self.relu)
        self.wpack = torch.cat([pack_umma_b(Wt) for Wt, _ in layers]).contiguous()
        self.bias = torch.cat([b.float() for _, b in layers]).contiguous()


def pack_umma_b(Wt):
    """[K,N] fp32 (y = x @ Wt) -> uint8 image of the tcgen05 B operand: fp16 hi then fp16 lo, each K/64 blocks of
    [N rows x 64 input channels] with 128-byte rows and the 128B swizzle (16-byte chunk c of row r at c ^ (r & 7))."""
    K, N = Wt.shape
    W = Wt.t().contiguous().float()                 # [N, K]: row n = output channel, K contiguous (K-major)
    hi = W.half()
    lo = (W - hi.float()).half()
    r7 = (torch.arange(N, device=W.device) & 7)[:, None]
    pos = torch.arange(8, device=W.device)[None, :] ^ r7           # chunk stored at position p holds source chunk p ^ (r & 7)
    idx = pos[:, None, :, None].expand(N, K // 64, 8, 8)
    out = []
    for h in (hi, lo):
        blocks = h.reshape(N, K // 64, 8, 8).gather(2, idx)        # [n, kc, position, element]
        out.append(blocks.permute(1, 0, 2, 3).contiguous().view(torch.uint8).reshape(-1))
    return torch.cat(out)


def mlp_chain(x, chain, add_in=None, add_out=None, want_y=True, want_packed=False):
    """y = chain(x + add_in) + add_out; returns (y fp32 or None, packed fp16 [rows, split_kp(N)] or None)."""
    lib = load()
    rows, K = x.shape
    assert K == chain.dims[0]
    N = chain.dims[-1]
    y = torch.empty(rows, N, dtype=torch.float32, device=x.device) if want_y else None
    Kp = split_kp(N)
    yp = torch.empty(rows, Kp, dtype=torch.float16, device=x.device) if want_packed else None
    rc = lib.pdsc_mlp_chain(_ptr(x, torch.float32, "x"), _ptr(add_in, torch.float32, "add_in"), rows, len(chain.relu),
                            ctypes.cast(chain.c_dims, c_void_p), ctypes.cast(chain.c_relu, c_void_p),
                            _ptr(chain.wpack, torch.uint8, "wpack"), _ptr(chain.bias, torch.float32, "bias"),
                            _ptr(add_out, torch.float32, "add_out"), _ptr(y), _ptr(yp), Kp, _stream())
    _check(rc, "pdsc_mlp_chain")
    _count()
    return y, yp


def mlp_chain_direct(x, chain, residual=None, residual_layer=0):
    """y = chain(x) with `residual` added to the output of layer `residual_layer` (0-based) before the next layer; up to four
    layers, rows straight between global memory and registers (pdsc_mlp_chain_direct)."""
    lib = load()
    rows, K = x.shape
    assert K == chain.dims[0]
    if residual is not None:
        assert residual.shape == (rows, chain.dims[residual_layer + 1])
    y = torch.empty(rows, chain.dims[-1], dtype=torch.float32, device=x.device)
    rc = lib.pdsc_mlp_chain_direct(_ptr(x, torch.float32, "x"), rows, len(chain.relu), ctypes.cast(chain.c_dims, c_void_p),
                                   ctypes.cast(chain.c_relu, c_void_p), _ptr(chain.wpack, torch.uint8, "wpack"),
                                   _ptr(chain.bias, torch.float32, "bias"), _ptr(residual, torch.float32, "residual"),
                                   int(residual_layer), _ptr(y), _stream())
    _check(rc, "pdsc_mlp_chain_direct")
    _count()
    return y


def pick_seeds(src, scores, R, S, ws=None, want_aux=False):
    lib = load()
    B, N = scores.shape
    seeds = torch.empty(B, S, dtype=torch.int64, device=scores.device)
    lm = torch.empty(B, N, dtype=torch.uint8, device=scores.device) if want_aux else None
    npos = torch.empty(B, dtype=torch.int32, device=scores.device) if want_aux else None
    w = _ws(workspace_bytes(OP_PICK_SEEDS, B, N), scores.device, ws)
    rc = lib.pdsc_pick_seeds(_ptr(src, torch.float32, "src"), _ptr(scores, torch.float32, "scores"), B, N, float(R), S,
                             seeds.data_ptr(), _ptr(lm), _ptr(npos), w.data_ptr(), w.numel(), _stream())
    _check(rc, "pdsc_pick_seeds")
    _count()
    return (seeds, lm, npos) if want_aux else seeds


def seed_knn(normed, seeds, k, ws=None, want_dist=False):
    lib = load()
    B, N, C = normed.shape
    S = seeds.shape[1]
    idx = torch.empty(B, S, k, dtype=torch.int64, device=normed.device)
    dist = torch.empty(B, S, k, dtype=torch.float32, device=normed.device) if want_dist else None
    w = _ws(workspace_bytes(OP_SEED_KNN, B, N, C, S, k), normed.device, ws)
    rc = lib.pdsc_seed_knn(_ptr(normed, torch.float32, "normed"), _ptr(seeds, torch.int64, "seeds"), B, N, C, S, k,
                           idx.data_ptr(), _ptr(dist), w.data_ptr(), w.numel(), _stream())
    _check(rc, "pdsc_seed_knn")
    _count()
    return (idx, dist) if want_dist else idx


def seed_hypotheses(normed, src, tgt, knn_idx, sigma, sigma_spat, num_iterations, ws=None, want_aux=False):
    lib = load()
    B, N, C = normed.shape
    S, k = knn_idx.shape[1], knn_idx.shape[2]
    trans = torch.empty(B, S, 4, 4, dtype=torch.float32, device=normed.device)
    weights = torch.empty(B, S, k, dtype=torch.float32, device=normed.device) if want_aux else None
    iters = torch.empty(B, dtype=torch.int32, device=normed.device) if want_aux else None
    w = _ws(workspace_bytes(OP_SEED_HYPOTHESES, B, N, C, S, k), normed.device, ws)
    rc = lib.pdsc_seed_hypotheses(_ptr(normed, torch.float32, "normed"), _ptr(src, torch.float32, "src"),
                                  _ptr(tgt, torch.float32, "tgt"), _ptr(knn_idx, torch.int64, "knn_idx"),
                                  B, N, C, S, k, float(sigma), float(sigma_spat), int(num_iterations),
                                  trans.data_ptr(), _ptr(weights), _ptr(iters), w.data_ptr(), w.numel(), _stream())
    _check(rc, "pdsc_seed_hypotheses")
    _count()
    return (trans, weights, iters) if want_aux else trans


def score_hypotheses(seed_trans, src, tgt, thr):
    lib = load()
    B, S = seed_trans.shape[0], seed_trans.shape[1]
    N = src.shape[1]
    dev = src.device
    counts = torch.empty(B, S, dtype=torch.int32, device=dev)
    best = torch.empty(B, dtype=torch.int64, device=dev)
    final_trans = torch.empty(B, 4, 4, dtype=torch.float32, device=dev)
    labels = torch.empty(B, N, dtype=torch.float32, device=dev)
    rc = lib.pdsc_score_hypotheses(_ptr(seed_trans, torch.float32, "seed_trans"), _ptr(src, torch.float32, "src"),
                                   _ptr(tgt, torch.float32, "tgt"), B, N, S, float(thr), counts.data_ptr(),
                                   best.data_ptr(), final_trans.data_ptr(), labels.data_ptr(), None, 0, _stream())
    _check(rc, "pdsc_score_hypotheses")
    _count()
    return counts, best, final_trans, labels


def post_refine(trans, src, tgt, tau, max_iter=20, want_aux=False):
    """Refines ``trans`` [B,4,4] IN PLACE and returns it."""
    lib = load()
    B, N = src.shape[0], src.shape[1]
    iters = torch.empty(B, dtype=torch.int32, device=src.device) if want_aux else None
    counts = torch.empty(B, max_iter, dtype=torch.int32, device=src.device) if want_aux else None
    rc = lib.pdsc_post_refine(_ptr(trans, torch.float32, "trans"), _ptr(src, torch.float32, "src"),
                              _ptr(tgt, torch.float32, "tgt"), B, N, float(tau), int(max_iter), _ptr(iters),
                              _ptr(counts), _stream())
    _check(rc, "pdsc_post_refine")
    _count()
    return (trans, iters, counts) if want_aux else trans


def feature_head(feat, w1t, b1, w2t, b2, w3, b3, eps=1e-12):
    """feat [B,N,C] -> (F.normalize(feat, dim=-1) [B,N,C], classifier logits [B,N]); models/PointDSC.py:156,171."""
    lib = load()
    B, N, C = feat.shape
    normed = torch.empty_like(feat)
    conf = torch.empty(B, N, dtype=torch.float32, device=feat.device)
    rc = lib.pdsc_feature_head(_ptr(feat, torch.float32, "feat"), B * N, C, int(w1t.shape[1]), int(w2t.shape[1]),
                               _ptr(w1t, torch.float32, "w1t"), _ptr(b1, torch.float32, "b1"), _ptr(w2t, torch.float32, "w2t"),
                               _ptr(b2, torch.float32, "b2"), _ptr(w3, torch.float32, "w3"), _ptr(b3, torch.float32, "b3"),
                               float(eps), normed.data_ptr(), conf.data_ptr(), _stream())
    _check(rc, "pdsc_feature_head")
    _count()
    return normed, conf


def compat_matvec(src, tgt, v, row_begin, row_end, sigma_spat, converged_flag=None, out=None):
    """y = M[row_begin:row_end, :] v for one pair; src,tgt [N,3], v [N]."""
    lib = load()
    N = v.shape[0]
    if out is None:
        out = torch.empty(row_end - row_begin, dtype=torch.float32, device=v.device)
    rc = lib.pdsc_compat_matvec(_ptr(src, torch.float32, "src"), _ptr(tgt, torch.float32, "tgt"),
                                _ptr(v, torch.float32, "v"), _ptr(out, torch.float32, "y"), N, int(row_begin),
                                int(row_end), float(sigma_spat), _ptr(converged_flag), _stream())
    _check(rc, "pdsc_compat_matvec")
    _count()
    return out


def power_normalize(y, v, converged_flag, iters_done=None):
    lib = load()
    rc = lib.pdsc_power_normalize(_ptr(y, torch.float32, "y"), _ptr(v, torch.float32, "v"), v.shape[0],
                                  _ptr(converged_flag, torch.int32, "flag"), _ptr(iters_done), None, 0, _stream())
    _check(rc, "pdsc_power_normalize")
    _count()
    return v


def power_iter_full(src, tgt, sigma_spat, num_iterations, ws=None):
    """Leading eigenvector of the zero-diagonal compatibility matrix of one pair ([N,3] inputs)."""
    lib = load()
    N = src.shape[0]
    v = torch.empty(N, dtype=torch.float32, device=src.device)
    iters = torch.empty(1, dtype=torch.int32, device=src.device)
    w = _ws(workspace_bytes(OP_POWER_ITER_FULL, 1, N), src.device, ws)
    rc = lib.pdsc_power_iter_full(_ptr(src, torch.float32, "src"), _ptr(tgt, torch.float32, "tgt"), v.data_ptr(), N,
                                  float(sigma_spat), int(num_iterations), iters.data_ptr(), w.data_ptr(), w.numel(),
                                  _stream())
    _check(rc, "pdsc_power_iter_full")
    _count()
    return v, iters


def set_overflow_flag(flag):
    """Registers (for this host thread) the int32 device tensor into which the split-fp16 kernels OR 1 when an operand is
    outside the fp16 range; None unregisters."""
    _check(load().pdsc_set_overflow_flag(None if flag is None else _ptr(flag, torch.int32, "overflow flag")), "pdsc_set_overflow_flag")


POWER_ITER_STATIC_MAX_N = 65536
P2P_MAX_WORLD = 8


def power_p2p_exchange_bytes(N):
    return int(load().pdsc_power_p2p_exchange_bytes(int(N)))


def compat_matvec_p2p(src, tgt, v, row_begin, row_end, sigma_spat, converged_flag, peer_ptrs, rank, seq):
    """One rank's rows of M v, stored into every rank's exchange buffer (peer_ptrs: device addresses, as ints)."""
    lib = load()
    world = len(peer_ptrs)
    arr = (c_void_p * world)(*[int(p) for p in peer_ptrs])
    rc = lib.pdsc_compat_matvec_p2p(_ptr(src, torch.float32, "src"), _ptr(tgt, torch.float32, "tgt"), _ptr(v, torch.float32, "v"),
                                    v.shape[0], int(row_begin), int(row_end), float(sigma_spat), _ptr(converged_flag),
                                    ctypes.cast(arr, c_void_p), world, int(rank), int(seq), _stream())
    _check(rc, "pdsc_compat_matvec_p2p")
    _count()


def power_normalize_p2p(exchange_ptr, v, converged_flag, iters_done, world, seq):
    lib = load()
    rc = lib.pdsc_power_normalize_p2p(int(exchange_ptr), _ptr(v, torch.float32, "v"), v.shape[0],
                                      _ptr(converged_flag, torch.int32, "flag"), _ptr(iters_done), int(world), int(seq), _stream())
    _check(rc, "pdsc_power_normalize_p2p")
    _count()
    return v


def power_iter_full_static(src, tgt, sigma_spat, num_iterations, v_out, iters_out=None):
    """pdsc_power_iter_full_static: as power_iter_full, but the library's own static device scratch is used and NOTHING
    is allocated here either -- the caller passes the outputs (v_out [N] fp32, iters_out [1] int32 or None).
    N <= POWER_ITER_STATIC_MAX_N; calls that overlap on the GPU share the scratch and race."""
    lib = load()
    N = src.shape[0]
    rc = lib.pdsc_power_iter_full_static(_ptr(src, torch.float32, "src"), _ptr(tgt, torch.float32, "tgt"),
                                         _ptr(v_out, torch.float32, "v_out"), N, float(sigma_spat), int(num_iterations),
                                         _ptr(iters_out, torch.int32, "iters_out"), _stream())
    _check(rc, "pdsc_power_iter_full_static")
    _count()
    return v_out


MATCH_DOT, MATCH_L2 = 0, 1


def nn_match(A, B, metric=MATCH_L2, want_second=False, want_values=False):
    """A [P,Na,D], B [P,Nb,D] fp32 -> idx1 int64 [P,Na] (nearest B row per A row, ties to the lowest index), plus
    idx2 (second nearest) and the two distances when asked for (pdsc_nn_match)."""
    lib = load()
    P, Na, D = A.shape
    Nb = B.shape[1]
    dev = A.device
    idx1 = torch.empty(P, Na, dtype=torch.int64, device=dev)
    idx2 = torch.empty(P, Na, dtype=torch.int64, device=dev) if want_second else None
    v1 = torch.empty(P, Na, dtype=torch.float32, device=dev) if want_values else None
    v2 = torch.empty(P, Na, dtype=torch.float32, device=dev) if want_values else None
    rc = lib.pdsc_nn_match(_ptr(A, torch.float32, "A"), _ptr(B, torch.float32, "B"), P, Na, Nb, D, metric, idx1.data_ptr(),
                           idx2.data_ptr() if idx2 is not None else None, v1.data_ptr() if v1 is not None else None,
                           v2.data_ptr() if v2 is not None else None, _stream())
    _check(rc, "pdsc_nn_match")
    _count()
    return idx1, idx2, v1, v2


def mutual_corr(idx_ab, idx_ba, Nb, use_mutual=True):
    """idx_ab int64 [P,Na], idx_ba int64 [P,Nb] -> (corr int64 [P,Na,2], count int32 [P]); rows [0,count) are valid."""
    lib = load()
    P, Na = idx_ab.shape
    corr = torch.empty(P, Na, 2, dtype=torch.int64, device=idx_ab.device)
    count = torch.empty(P, dtype=torch.int32, device=idx_ab.device)
    rc = lib.pdsc_mutual_corr(_ptr(idx_ab, torch.int64, "idx_ab"), _ptr(idx_ba, torch.int64, "idx_ba") if idx_ba is not None else None,
                              P, Na, Nb, 1 if use_mutual else 0, corr.data_ptr(), count.data_ptr(), _stream())
    _check(rc, "pdsc_mutual_corr")
    _count()
    return corr, count


def registration_metrics(trans, gt_trans, src, tgt):
    """[P,4,4] x2, [P,N,3] x2 -> fp32 [P,4] = RE [deg], TE [cm], mean residual, mean squared residual."""
    lib = load()
    P, N, _ = src.shape
    out = torch.empty(P, 4, dtype=torch.float32, device=src.device)
    rc = lib.pdsc_registration_metrics(_ptr(trans, torch.float32, "trans"), _ptr(gt_trans, torch.float32, "gt_trans"),
                                       _ptr(src, torch.float32, "src"), _ptr(tgt, torch.float32, "tgt"), P, N, out.data_ptr(), _stream())
    _check(rc, "pdsc_registration_metrics")
    _count()
    return out


def classification_counts(pred, gt, thr=0.0):
    """pred, gt fp32 [P,N] -> (int32 [P,4] tp fp fn tn, fp32 [P,2] sum pred*gt, sum pred*(1-gt))."""
    lib = load()
    P, N = pred.shape
    counts = torch.empty(P, 4, dtype=torch.int32, device=pred.device)
    sums = torch.empty(P, 2, dtype=torch.float32, device=pred.device)
    rc = lib.pdsc_classification_counts(_ptr(pred, torch.float32, "pred"), _ptr(gt, torch.float32, "gt"), P, N, float(thr),
                                        counts.data_ptr(), sums.data_ptr(), _stream())
    _check(rc, "pdsc_classification_counts")
    _count()
    return counts, sums


# ---- (N1) demo front end -------------------------------------------------------------------------
def hybrid_search(pts, radius, max_nn, want_lists=True, want_normals=False, ws=None):
    """pts fp64 [n,3] -> (idx int32 [n,max_nn], dist2 fp64 [n,max_nn], count int32 [n], normals fp64 [n,3] or None)
    (pdsc_hybrid_search).  With want_lists=False only the normals are produced."""
    lib = load()
    n = pts.shape[0]
    dev = pts.device
    idx = torch.empty(n, max_nn, dtype=torch.int32, device=dev) if want_lists else None
    d2 = torch.empty(n, max_nn, dtype=torch.float64, device=dev) if want_lists else None
    cnt = torch.empty(n, dtype=torch.int32, device=dev)
    nrm = torch.empty(n, 3, dtype=torch.float64, device=dev) if want_normals else None
    buf = _ws(workspace_bytes(OP_HYBRID_SEARCH, N=n), dev, ws)
    rc = lib.pdsc_hybrid_search(_ptr(pts, torch.float64, "pts"), n, float(radius), int(max_nn),
                                idx.data_ptr() if want_lists else None, d2.data_ptr() if want_lists else None, cnt.data_ptr(),
                                nrm.data_ptr() if want_normals else None, buf.data_ptr(), buf.numel(), _stream())
    _check(rc, "pdsc_hybrid_search")
    _count()
    return idx, d2, cnt, nrm


def voxel_down_sample(pts, normals, voxel, cells=1 << 22):
    """pts fp64 [n,3], normals fp64 [n,3] or None -> (points [m,3], normals [m,3] or None), one row per occupied voxel
    in ascending (ix,iy,iz) (pdsc_voxel_down_sample).  Reads the row count back (one host sync)."""
    lib = load()
    n = pts.shape[0]
    dev = pts.device
    out_p = torch.empty(n, 3, dtype=torch.float64, device=dev)
    out_n = torch.empty(n, 3, dtype=torch.float64, device=dev) if normals is not None else None
    count = torch.empty(1, dtype=torch.int32, device=dev)
    for _ in range(2):
        buf = torch.empty(workspace_bytes(OP_VOXEL_DOWN_SAMPLE, N=n, S=cells), dtype=torch.uint8, device=dev)
        rc = lib.pdsc_voxel_down_sample(_ptr(pts, torch.float64, "pts"),
                                        _ptr(normals, torch.float64, "normals") if normals is not None else None, n,
                                        float(voxel), out_p.data_ptr(), out_n.data_ptr() if out_n is not None else None,
                                        count.data_ptr(), buf.data_ptr(), buf.numel(), _stream())
        _check(rc, "pdsc_voxel_down_sample")
        _count()
        m = int(count.item())
        if m >= 0:
            return out_p[:m], (out_n[:m] if out_n is not None else None)
        cells = int(buf[56:60].view(torch.int32).item()) + 1024      # cells the grid needs (see the header)
        if cells > (1 << 28):
            break
    raise RuntimeError("pdsc_voxel_down_sample: the voxel grid over the bounding box is too large for a dense grid")


def fpfh(pts, normals, idx, d2, cnt):
    """FPFH-33 fp64 [n,33] from the hybrid-search lists of the same cloud (pdsc_fpfh)."""
    lib = load()
    n, max_nn = idx.shape
    spfh = torch.empty(n, 33, dtype=torch.float64, device=pts.device)
    out = torch.empty(n, 33, dtype=torch.float64, device=pts.device)
    rc = lib.pdsc_fpfh(_ptr(pts, torch.float64, "pts"), _ptr(normals, torch.float64, "normals"), n, _ptr(idx, torch.int32, "idx"),
                       _ptr(d2, torch.float64, "dist2"), _ptr(cnt, torch.int32, "count"), max_nn, spfh.data_ptr(), out.data_ptr(),
                       _stream())
    _check(rc, "pdsc_fpfh")
    _count()
    return out


def nn_match_f64(A, B, want_values=False, ws=None):
    """A fp64 [Na,D], B fp64 [Nb,D] -> idx int64 [Na] = argmin_j sqrt(2 - 2 a.b + 1e-6) (pdsc_nn_match_f64)."""
    lib = load()
    Na, D = A.shape
    idx = torch.empty(Na, dtype=torch.int64, device=A.device)
    val = torch.empty(Na, dtype=torch.float64, device=A.device) if want_values else None
    buf = _ws(workspace_bytes(OP_NN_MATCH_F64, N=Na), A.device, ws)
    rc = lib.pdsc_nn_match_f64(_ptr(A, torch.float64, "A"), _ptr(B, torch.float64, "B"), Na, B.shape[0], D, idx.data_ptr(),
                               val.data_ptr() if want_values else None, buf.data_ptr(), buf.numel(), _stream())
    _check(rc, "pdsc_nn_match_f64")
    _count()
    return (idx, val) if want_values else idx


# ---- (N4) ICP hand-off ---------------------------------------------------------------------------
def icp_refine(src, tgt, trans, max_dist=0.1, max_iter=30, rel_fitness=1e-6, rel_rmse=1e-6, want_corr=False, ws=None):
    """src, tgt fp32 [B,N,3], trans fp32 [B,4,4] -> (refined trans fp32 [B,4,4], stats fp64 [B,2] = fitness / inlier rmse,
    iters int32 [B], corr int32 [B,N] or None) (pdsc_icp_refine)."""
    lib = load()
    B, N, _ = src.shape
    dev = src.device
    out = trans.detach().clone().contiguous()
    stats = torch.empty(B, 2, dtype=torch.float64, device=dev)
    iters = torch.empty(B, dtype=torch.int32, device=dev)
    corr = torch.empty(B, N, dtype=torch.int32, device=dev) if want_corr else None
    buf = _ws(workspace_bytes(OP_ICP_REFINE, B=B, N=N), dev, ws)
    rc = lib.pdsc_icp_refine(_ptr(src, torch.float32, "src"), _ptr(tgt, torch.float32, "tgt"), _ptr(out, torch.float32, "trans"), B, N,
                             float(max_dist), int(max_iter), float(rel_fitness), float(rel_rmse), stats.data_ptr(), iters.data_ptr(),
                             corr.data_ptr() if want_corr else None, buf.data_ptr(), buf.numel(), _stream())
    _check(rc, "pdsc_icp_refine")
    _count()
    return out, stats, iters, corr
