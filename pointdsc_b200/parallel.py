"""Multi-GPU pieces of the hot path (one process per GPU, torch.distributed).

* Registration pairs are independent: ``shard_pairs`` gives each rank its slice, no data-path
  collective (the pattern of the reference's test_parallel.sh:18-24 / DistributedSampler).
* One very large pair (N >= 50k, BASELINE config 4): the implicit N x N compatibility matrix is
  row-sharded.  Every power iteration each rank computes y[r0:r1] = M[r0:r1,:] v and every rank
  normalises redundantly -- which also evaluates the reference's allclose stopping rule on the
  device, so no second exchange and no host sync is needed (cal_leading_eigenvector,
  models/PointDSC.py:338-358).  Two transports for the 4N bytes of y:
  ``power_iteration_full_p2p``      the product path on one NVSwitch box: the matvec kernel stores its rows into every
                                    rank's buffer over NVLink and publishes an arrival flag, the normalise kernel waits
                                    for the flags on the device (``PeerExchange``: symmetric memory; no collective);
  ``power_iteration_full_sharded``  one all-gather per iteration (NCCL on GPUs, gloo in the CPU tests): the baseline
                                    transport, also what runs when peer mappings are unavailable.
"""
from __future__ import annotations

import torch


def shard_range(n: int, rank: int, world: int):
    """Contiguous, balanced [begin, end) of n items for `rank` (first n % world ranks get one more)."""
    base, extra = divmod(n, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def shard_pairs(n_pairs: int, rank: int, world: int):
    return list(range(*shard_range(n_pairs, rank, world)))


def _default_ops():
    from . import _cabi
    return _cabi.compat_matvec, _cabi.power_normalize


def power_iteration_full_sharded(src, tgt, sigma_spat, num_iterations, group=None, ops=None):
    """Leading eigenvector of the zero-diagonal compatibility matrix of ONE pair, rows sharded over
    the ranks of `group`.  src,tgt: [N,3] (replicated on every rank).  Returns (v [N], iterations
    used, as a 1-element int32 tensor).  `ops` = (matvec, normalize) is injectable so the host
    logic can be exercised on CPU with gloo in the tests; the product path uses the CUDA kernels."""
    import torch.distributed as dist
    matvec, normalize = ops or _default_ops()
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    N = src.shape[0]
    r0, r1 = shard_range(N, rank, world)
    v = torch.ones(N, dtype=torch.float32, device=src.device)
    flag = torch.zeros(1, dtype=torch.int32, device=src.device)
    iters = torch.zeros(1, dtype=torch.int32, device=src.device)
    # all_gather needs equal shard sizes: pad every shard to the largest one
    width = (N + world - 1) // world
    y_full = torch.empty(world * width, dtype=torch.float32, device=src.device)
    y_loc = torch.zeros(width, dtype=torch.float32, device=src.device)
    bounds = [shard_range(N, r, world) for r in range(world)]
    y = torch.empty(N, dtype=torch.float32, device=src.device)
    for _ in range(num_iterations):
        matvec(src, tgt, v, r0, r1, sigma_spat, converged_flag=flag, out=y_loc[: r1 - r0])
        if world > 1:
            dist.all_gather_into_tensor(y_full, y_loc, group=group)
            if world * width == N:                      # equal shards: the gathered buffer is y
                y = y_full
            else:
                for r, (a, b) in enumerate(bounds):
                    y[a:b] = y_full[r * width: r * width + (b - a)]
        else:
            y[:] = y_loc[:N]
        normalize(y, v, flag, iters)
    return v, iters


class PeerExchange:
    """Symmetric-memory exchange buffer of the row-sharded power iteration (one per rank; ``pdsc_compat_matvec_p2p`` /
    ``pdsc_power_normalize_p2p`` in include/pointdsc_b200.h).  torch's symmetric memory supplies what the C ABI needs
    and nothing else: a buffer of the same size on every rank and, on each rank, the device addresses at which it sees
    its peers' buffers (NVLink peer mappings).  No NCCL call is made per iteration."""

    def __init__(self, N, device, group=None):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        from . import _cabi
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        if self.world > _cabi.P2P_MAX_WORLD:
            raise RuntimeError(f"P2P exchange supports up to {_cabi.P2P_MAX_WORLD} ranks (one NVSwitch box), got {self.world}")
        self.N = int(N)
        nbytes = _cabi.power_p2p_exchange_bytes(self.N)
        self.buf = symm_mem.empty(nbytes, dtype=torch.uint8, device=device)
        self.handle = symm_mem.rendezvous(self.buf, self.group)
        self.peer_ptrs = [int(p) for p in self.handle.buffer_ptrs]
        assert self.peer_ptrs[self.rank] == self.buf.data_ptr()
        self.buf.zero_()                      # flags and counter start at zero
        torch.cuda.synchronize(device)
        dist.barrier(self.group)              # ... on every rank, before anybody stores into a peer
        self.seq = 0                          # iterations issued so far (identical on all ranks)


def power_iteration_full_p2p(src, tgt, sigma_spat, num_iterations, exchange):
    """Row-sharded leading eigenvector of ONE pair with the eigenvector exchanged by direct NVLink stores (see
    PeerExchange): per iteration one matvec launch (rows of this rank -> every rank's y) and one normalise launch (waits
    on the device for every rank's arrival flag).  Same results, bit for bit, as one GPU.  Returns (v [N], iterations)."""
    from . import _cabi
    N = src.shape[0]
    if N != exchange.N:
        raise ValueError(f"exchange buffer was made for N={exchange.N}, got N={N}")
    r0, r1 = shard_range(N, exchange.rank, exchange.world)
    v = torch.ones(N, dtype=torch.float32, device=src.device)
    flag = torch.zeros(1, dtype=torch.int32, device=src.device)
    iters = torch.zeros(1, dtype=torch.int32, device=src.device)
    for _ in range(num_iterations):
        exchange.seq += 1
        _cabi.compat_matvec_p2p(src, tgt, v, r0, r1, sigma_spat, flag, exchange.peer_ptrs, exchange.rank, exchange.seq)
        _cabi.power_normalize_p2p(exchange.peer_ptrs[exchange.rank], v, flag, iters, exchange.world, exchange.seq)
    return v, iters
