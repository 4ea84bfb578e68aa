// (b) Power iteration on the full N x N spatial-compatibility matrix (zero diagonal):
// cal_leading_eigenvector models/PointDSC.py:338-358 on the matrix of :151-153, the
// spectral-matching use kept at models/PointDSC.py:170 and baseline_scripts/baseline_3DMatch.py:19-46.
// The matrix (10 GB at N=50k) is never stored: each M.v recomputes it from 6 floats per
// correspondence.  Rows are independent, so a rank can own a row range (row sharding) and
// all-gather y; the normalisation + allclose test is a one-block kernel every rank repeats.
#include <cooperative_groups.h>
#include <math.h>

#include "common.cuh"
#include "sm100.cuh"

namespace cg = cooperative_groups;

namespace pdsc {

// 64 rows per cluster (32 lanes x 2 rows per thread), 8 column slices per CTA (one warp each: every lane of a warp reads the
// same column, a shared-memory broadcast).  Round 1 used 128 rows x 4 slices: a rank that owns N/8 = 6250 rows then launched
// 392 CTAs, a third of one wave on 148 SMs; with 64-row blocks it launches 784 and the single-GPU run 6256.
constexpr int MV_LANES = 32, MV_RPT = 2, MV_ROWS = MV_LANES * MV_RPT, MV_SLICES = 8, MV_THREADS = MV_LANES * MV_SLICES,
              MV_TILE = 512;
constexpr int MV_KS = 8;        // most column splits = CTAs per cluster (the launch picks 4..8, see mv_column_splits)

__device__ __forceinline__ float mv_fma_sat(float a, float b, float c) {
    float d;
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

// y[r] = sum_j M[r,j] v[j], M[r,j] = max(0, 1 - (|s_r - s_j| - |t_r - t_j|)^2 / sigma^2), M[r,r] = 0, for MV_ROWS (64) rows per
// CLUSTER of 8 CTAs: CTA k of the cluster takes the k-th eighth of the column tiles, the eight partial sums meet in
// distributed shared memory and rank 0 adds them in a fixed order.  The split does not depend on the row range, so a
// row-sharded run (any number of ranks) produces the same bits as a single GPU, and a rank that owns only N/8 rows still
// fills the machine.  Arithmetic: points are scaled by 1/sigma when a tile is staged (no division per entry), a thread
// owns two rows as the two lanes of packed f32x2 adds / FMAs, distances by sqrt.approx (MUFU), 1 - c^2 by one saturating
// FMA: 2 MUFU + ~9 packed FP32 instructions per pair of entries (the exact-rounding form it replaces cost ~50 per entry).
//
// P2P = true (row-sharded run over NVLink, config 4): rank 0 of every cluster stores its 64 results straight into the y
// buffer of EVERY rank (peer device pointers into symmetric memory; y is indexed by the global row, so there is no
// padding and no gather step), and the last cluster of the grid to finish (device-scope counter) publishes this rank's
// arrival: a system-scope release store of `seq` into flag slot `rank` of every peer.  The normalise kernel of each rank
// waits for all `world` slots to reach `seq` (power_normalize_kernel), so one iteration is two launches and no collective.
struct MvPeers {
    float* y[PDSC_P2P_MAX_WORLD];          // y buffer (of this iteration's parity) on every rank
    int32_t* flags[PDSC_P2P_MAX_WORLD];    // arrival flags [world] on every rank
    int32_t* done;                          // this rank's cluster counter
    int world, rank, seq;
};

__device__ __forceinline__ void st_release_sys(int32_t* p, int32_t v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int32_t ld_acquire_sys(const int32_t* p) {
    int32_t v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

template <bool P2P>
__global__ void __launch_bounds__(MV_THREADS)
compat_matvec_kernel(const float* __restrict__ src, const float* __restrict__ tgt, const float* __restrict__ v,
                     float* __restrict__ y, int N, int row_begin, int row_end, float inv_sigma,
                     const int32_t* __restrict__ converged_flag, const MvPeers peers) {
    if (converged_flag && *converged_flag) return;          // uniform over the grid (and over the ranks: every rank normalises the same y)
    cg::cluster_group cluster = cg::this_cluster();
    __shared__ float4 ta[MV_TILE];   // src xyz / sigma, v
    __shared__ float4 tb[MV_TILE];   // tgt xyz / sigma, unused
    __shared__ float2 part[MV_SLICES][MV_LANES];
    __shared__ float2 csum[MV_LANES];  // this CTA's partial sums (rows rl and rl + 64), read by rank 0 of the cluster
    const int rl = threadIdx.x % MV_LANES, slice = threadIdx.x / MV_LANES;
    const int ks = blockIdx.y;                                // == rank in the cluster
    const int r0 = row_begin + blockIdx.x * MV_ROWS + rl, r1 = r0 + MV_LANES;
    const int c0 = min(r0, N - 1), c1 = min(r1, N - 1);
    const float2 sx = make_float2(src[c0 * 3] * inv_sigma, src[c1 * 3] * inv_sigma),
                 sy = make_float2(src[c0 * 3 + 1] * inv_sigma, src[c1 * 3 + 1] * inv_sigma),
                 sz = make_float2(src[c0 * 3 + 2] * inv_sigma, src[c1 * 3 + 2] * inv_sigma);
    const float2 tx = make_float2(tgt[c0 * 3] * inv_sigma, tgt[c1 * 3] * inv_sigma),
                 ty = make_float2(tgt[c0 * 3 + 1] * inv_sigma, tgt[c1 * 3 + 1] * inv_sigma),
                 tz = make_float2(tgt[c0 * 3 + 2] * inv_sigma, tgt[c1 * 3 + 2] * inv_sigma);
    const int ntiles = (N + MV_TILE - 1) / MV_TILE;
    const int KS = (int)gridDim.y;                            // column splits = cluster size (launch attribute)
    const int t_lo = (int)((long long)ntiles * ks / KS), t_hi = (int)((long long)ntiles * (ks + 1) / KS);
    float2 acc = make_float2(0.f, 0.f);
    for (int tile = t_lo; tile < t_hi; ++tile) {
        const int j0 = tile * MV_TILE;
        const int nj = min(MV_TILE, N - j0);
        __syncthreads();
        for (int t = threadIdx.x; t < nj; t += MV_THREADS) {
            const int j = j0 + t;
            ta[t] = make_float4(src[j * 3] * inv_sigma, src[j * 3 + 1] * inv_sigma, src[j * 3 + 2] * inv_sigma, v[j]);
            tb[t] = make_float4(tgt[j * 3] * inv_sigma, tgt[j * 3 + 1] * inv_sigma, tgt[j * 3 + 2] * inv_sigma, 0.f);
        }
        __syncthreads();
        // slice s takes elements s, s + MV_SLICES, ... of the tile: a fixed order, so y is reproducible
        const int d0 = r0 - j0, d1 = r1 - j0;                 // tile positions of the two diagonal entries (if inside)
#pragma unroll 4
        for (int t = slice; t < nj; t += MV_SLICES) {
            const float4 a = ta[t], c = tb[t];
            float2 dx = __fadd2_rn(sx, make_float2(-a.x, -a.x));
            float2 dy = __fadd2_rn(sy, make_float2(-a.y, -a.y));
            float2 dz = __fadd2_rn(sz, make_float2(-a.z, -a.z));
            float2 q = __ffma2_rn(dz, dz, __ffma2_rn(dy, dy, __fmul2_rn(dx, dx)));
            const float2 ds = make_float2(sm100::fast_sqrt(q.x), sm100::fast_sqrt(q.y));
            dx = __fadd2_rn(tx, make_float2(-c.x, -c.x));
            dy = __fadd2_rn(ty, make_float2(-c.y, -c.y));
            dz = __fadd2_rn(tz, make_float2(-c.z, -c.z));
            q = __ffma2_rn(dz, dz, __ffma2_rn(dy, dy, __fmul2_rn(dx, dx)));
            const float2 dt = make_float2(sm100::fast_sqrt(q.x), sm100::fast_sqrt(q.y));
            const float2 e = __fadd2_rn(ds, make_float2(-dt.x, -dt.y));
            float2 m = make_float2(mv_fma_sat(-e.x, e.x, 1.0f), mv_fma_sat(-e.y, e.y, 1.0f));     // max(0, 1 - e^2)
            if (t == d0) m.x = 0.f;
            if (t == d1) m.y = 0.f;
            acc = __ffma2_rn(m, make_float2(a.w, a.w), acc);
        }
    }
    part[slice][rl] = acc;
    __syncthreads();
    if (slice == 0) {
        static_assert(MV_SLICES == 8, "fixed-order tree below");
        const float2 p0 = part[0][rl], p1 = part[1][rl], p2 = part[2][rl], p3 = part[3][rl];
        const float2 p4 = part[4][rl], p5 = part[5][rl], p6 = part[6][rl], p7 = part[7][rl];
        csum[rl] = make_float2(((p0.x + p1.x) + (p2.x + p3.x)) + ((p4.x + p5.x) + (p6.x + p7.x)),
                               ((p0.y + p1.y) + (p2.y + p3.y)) + ((p4.y + p5.y) + (p6.y + p7.y)));
    }
    cluster.sync();
    if (ks == 0 && slice == 0) {
        float2 c[MV_KS];
#pragma unroll
        for (int k = 0; k < MV_KS; ++k) c[k] = k < KS ? cluster.map_shared_rank(csum, k)[rl] : make_float2(0.f, 0.f);
        float y0 = c[0].x, y1 = c[0].y;                       // fixed order (the same for every row range and world size)
#pragma unroll
        for (int k = 1; k < MV_KS; ++k) if (k < KS) { y0 += c[k].x; y1 += c[k].y; }
        if (!P2P) {
            if (r0 < row_end) y[r0 - row_begin] = y0;
            if (r1 < row_end) y[r1 - row_begin] = y1;
        } else {
            for (int p = 0; p < peers.world; ++p) {           // NVLink stores (coalesced 256-byte runs per peer)
                if (r0 < row_end) peers.y[p][r0] = y0;
                if (r1 < row_end) peers.y[p][r1] = y1;
            }
            __threadfence_system();                           // this thread's peer stores are ordered before the counter
        }
    }
    if (P2P && ks == 0) {
        __syncthreads();
        if (threadIdx.x == 0) {
            const int arrived = atomicAdd(peers.done, 1);
            if (arrived == (int)gridDim.x - 1) {              // last cluster of this rank: everything is written
                *peers.done = 0;                              // ready for the next launch
                __threadfence_system();
                for (int p = 0; p < peers.world; ++p) st_release_sys(peers.flags[p] + peers.rank, peers.seq);
            }
        }
    }
    cluster.sync();                                           // peers keep their shared memory until rank 0 has read it
}

// v_new = y / (|y| + 1e-6), the allclose test against v_last, the iteration counter: ONE cluster of PN_CTAS CTAs (the kernel
// every rank repeats on the whole y each iteration, so its latency is not divided by the number of GPUs: one CTA took ~45 us
// for N=50000, two passes of 49 dependent loads per thread).  Every CTA sums the squares of its slice in double, the
// partial sums meet in distributed shared memory and every CTA adds all of them in the same fixed order, so all CTAs (and
// all ranks) hold the same norm; the "all close" votes are combined the same way.
// wait_flags != nullptr (P2P run): y is complete once every rank's arrival flag has reached `seq`; the waiting threads poll
// with system-scope acquire loads (bounded: a lost peer traps instead of hanging the GPU), the barrier orders every other
// thread's reads of y behind them, and y is read past the L1 (peers wrote it over NVLink into this GPU's memory).
constexpr int PN_CTAS = 8, PN_THREADS = 1024;
__global__ void __cluster_dims__(PN_CTAS, 1, 1) __launch_bounds__(PN_THREADS)
power_normalize_kernel(const float* __restrict__ y, float* __restrict__ v, int N, int32_t* __restrict__ converged_flag,
                       int32_t* __restrict__ iters_done, const int32_t* __restrict__ wait_flags, int world, int seq) {
    if (*converged_flag) return;                             // uniform over the cluster (nobody has written it yet: see the end)
    cg::cluster_group cluster = cg::this_cluster();
    __shared__ double red[32];
    __shared__ int red_close[32];
    __shared__ double cta_sum;                               // read by the peers through DSMEM
    __shared__ int cta_close;
    if (wait_flags) {
        if ((int)threadIdx.x < world) {
            long long spins = 0;
            while (ld_acquire_sys(wait_flags + threadIdx.x) < seq) {
                __nanosleep(64);
                if (++spins > (1ll << 24)) __trap();          // ~2 s
            }
        }
        __syncthreads();
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rank = (int)cluster.block_rank();
    const int per = (N + PN_CTAS - 1) / PN_CTAS, i0 = rank * per, i1 = min(N, i0 + per);
    double ss = 0.0;
    for (int i = i0 + (int)threadIdx.x; i < i1; i += PN_THREADS) { const float yi = __ldcg(y + i); ss += (double)yi * (double)yi; }
    ss = warp_sum(ss);
    if (lane == 0) red[warp] = ss;
    __syncthreads();
    if (threadIdx.x == 0) {
        double tot = 0.0;
        for (int w = 0; w < 32; ++w) tot += red[w];
        cta_sum = tot;
    }
    cluster.sync();
    double tot = 0.0;
    for (int c = 0; c < PN_CTAS; ++c) tot += *cluster.map_shared_rank(&cta_sum, c);
    const float nrm = (float)sqrt(tot);
    int close = 1;
    for (int i = i0 + (int)threadIdx.x; i < i1; i += PN_THREADS) {
        const float vn = __fdiv_rn(__ldcg(y + i), nrm + 1e-6f);
        const float vl = v[i];
        close &= (fabsf(vn - vl) <= 1e-8f + 1e-5f * fabsf(vl)) ? 1 : 0;
        v[i] = vn;
    }
    close = __all_sync(0xffffffffu, close);
    if (lane == 0) red_close[warp] = close;
    __syncthreads();
    if (threadIdx.x == 0) {
        int all = 1;
        for (int w = 0; w < 32; ++w) all &= red_close[w];
        cta_close = all;
    }
    cluster.sync();
    if (rank == 0 && threadIdx.x == 0) {
        int all = 1;
        for (int c = 0; c < PN_CTAS; ++c) all &= *cluster.map_shared_rank(&cta_close, c);
        if (all) *converged_flag = 1;                        // (every CTA read the flag before the first cluster.sync)
        if (iters_done) *iters_done += 1;
    }
    cluster.sync();                                          // peers keep their shared memory until rank 0 has read it
}

__global__ void fill_ones_kernel(float* v, int N, int32_t* flag, int32_t* iters) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N) v[i] = 1.0f;
    if (i == 0) { *flag = 0; *iters = 0; }
}

// Column splits (= CTAs per cluster): 8 for every N and every row range -- never a function of the world size, so that a
// row-sharded run adds the same partial sums in the same order as one GPU.  Other splits were measured on the row range a
// rank owns at N = 50000 (scripts/prof_matvec.py, one B200, us per launch):
//   split   all rows   1/2      1/4      1/8
//     4      1627      862      447      236
//     6      1614      833      443      258
//     8      1624      826      446      235      <- 86 % of 1624 / 8 on 1/8 of the rows
// (a wave model -- 4 resident CTAs per SM, 6 x 98 clusters = one wave -- predicted 6; the 8-GPU run said 2.95 ms against
// 2.76 ms with 8.)  Non-portable cluster sizes, measured later on another box (8 there: 1645 / 848 / 455 / 243):
//    10      1591      808      416      244
//    12      1798      921      478      259
//    16      1863      950      497      264      <- larger clusters schedule worse; nothing beats 8 on 1/8 of the rows
// PDSC_MV_KS (tuning build) overrides it (up to 8).
static int mv_column_splits(int N) {
    (void)N;
    const int knob = tuning_int("PDSC_MV_KS", 0);
    return (knob >= 1 && knob <= MV_KS) ? knob : MV_KS;
}

template <bool P2P, typename... Args>
static void mv_launch(int nblk, int N, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    const int ks = mv_column_splits(N);
    cfg.gridDim = dim3(nblk, ks); cfg.blockDim = dim3(MV_THREADS); cfg.dynamicSmemBytes = 0; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 1; at[0].val.clusterDim.y = ks; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, compat_matvec_kernel<P2P>, args...);
}

int launch_compat_matvec(const float* src, const float* tgt, const float* v, float* y, int N, int row_begin,
                         int row_end, float sigma_spat, const int32_t* converged_flag, cudaStream_t st) {
    const int rows = row_end - row_begin;
    if (rows <= 0) return PDSC_OK;
    mv_launch<false>((rows + MV_ROWS - 1) / MV_ROWS, N, st, src, tgt, v, y, N, row_begin, row_end, 1.0f / sigma_spat, converged_flag, MvPeers{});
    count_launches(1);
    return check_launch("compat_matvec");
}

// Exchange buffer of the P2P run, one per rank, in memory every peer can address (symmetric memory):
//   [y parity 0: N floats | y parity 1: N floats | flags: PDSC_P2P_MAX_WORLD int32 | done counter: int32], 256-byte aligned parts
static size_t p2p_y_bytes(int N) { return align_up(sizeof(float) * (size_t)N, 256); }
size_t power_p2p_exchange_bytes(int N) { return 2 * p2p_y_bytes(N) + 256; }

int launch_compat_matvec_p2p(const float* src, const float* tgt, const float* v, int N, int row_begin, int row_end,
                             float sigma_spat, const int32_t* converged_flag, void* const* peer_bases, int world, int rank,
                             int seq, cudaStream_t st) {
    const int rows = row_end - row_begin;
    MvPeers peers;
    for (int p = 0; p < world; ++p) {
        char* base = reinterpret_cast<char*>(peer_bases[p]);
        peers.y[p] = reinterpret_cast<float*>(base + (size_t)(seq & 1) * p2p_y_bytes(N));
        peers.flags[p] = reinterpret_cast<int32_t*>(base + 2 * p2p_y_bytes(N));
    }
    peers.done = peers.flags[rank] + PDSC_P2P_MAX_WORLD;
    peers.world = world; peers.rank = rank; peers.seq = seq;
    // (a rank with no rows still has to publish its arrival: one cluster with an empty row range does it)
    const int nblk = rows > 0 ? (rows + MV_ROWS - 1) / MV_ROWS : 1;
    mv_launch<true>(nblk, N, st, src, tgt, v, (float*)nullptr, N, row_begin, row_end, 1.0f / sigma_spat, converged_flag, peers);
    count_launches(1);
    return check_launch("compat_matvec_p2p");
}

int launch_power_normalize_p2p(void* exchange_base, float* v, int N, int32_t* converged_flag, int32_t* iters_done, int world,
                               int seq, cudaStream_t st) {
    char* base = reinterpret_cast<char*>(exchange_base);
    const float* y = reinterpret_cast<const float*>(base + (size_t)(seq & 1) * p2p_y_bytes(N));
    const int32_t* flags = reinterpret_cast<const int32_t*>(base + 2 * p2p_y_bytes(N));
    power_normalize_kernel<<<PN_CTAS, PN_THREADS, 0, st>>>(y, v, N, converged_flag, iters_done, flags, world, seq);
    count_launches(1);
    return check_launch("power_normalize_p2p");
}

size_t power_workspace(int N) { return align_up(sizeof(float) * (size_t)N, 256) + 256; }

int launch_power_normalize(const float* y, float* v, int N, int32_t* converged_flag, int32_t* iters_done, void* ws,
                           size_t ws_bytes, cudaStream_t st) {
    (void)ws; (void)ws_bytes;
    power_normalize_kernel<<<PN_CTAS, PN_THREADS, 0, st>>>(y, v, N, converged_flag, iters_done, nullptr, 0, 0);
    count_launches(1);
    return check_launch("power_normalize");
}

int power_iter_full(const float* src, const float* tgt, float* v_out, int N, float sigma_spat, int num_iterations,
                    int32_t* iters_out, void* ws, size_t ws_bytes, cudaStream_t st) {
    if (ws_bytes < power_workspace(N)) { set_error("power_iter_full: workspace too small"); return PDSC_ERR_WORKSPACE; }
    float* y = reinterpret_cast<float*>(ws);
    int32_t* flag = reinterpret_cast<int32_t*>(reinterpret_cast<char*>(ws) + align_up(sizeof(float) * (size_t)N, 256));
    int32_t* iters = flag + 1;
    fill_ones_kernel<<<(N + 255) / 256, 256, 0, st>>>(v_out, N, flag, iters);
    count_launches(1);
    for (int it = 0; it < num_iterations; ++it) {
        int rc = launch_compat_matvec(src, tgt, v_out, y, N, 0, N, sigma_spat, flag, st);
        if (rc) return rc;
        rc = launch_power_normalize(y, v_out, N, flag, iters, nullptr, 0, st);
        if (rc) return rc;
    }
    if (iters_out) cudaMemcpyAsync(iters_out, iters, sizeof(int32_t), cudaMemcpyDeviceToDevice, st);
    return check_launch("power_iter_full");
}

// ------------------------------------------------------------------------------------------------
// No-allocation variant (pdsc_power_iter_full_static): the scratch of power_iter_full (y[N] + the converged flag + the
// iteration counter) lives in a file-scope device array instead of a caller-supplied workspace, for callers that cannot
// allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Every GPU that loads the library
// holds the array (256 KB + 256 B) until exit.  ALL calls on a device share it: calls that overlap on the GPU (different
// streams) race; calls on one stream are ordered and safe.  Nothing in it has to survive a call: fill_ones_kernel resets
// the flag and the counter at the start of every call and y is fully written before it is read.
static __device__ __align__(256) unsigned char g_power_scratch[sizeof(float) * PDSC_POWER_ITER_STATIC_MAX_N + 256];

// In host code `g_power_scratch` is only a stand-in for the symbol, not a device address: the address comes from
// cudaGetSymbolAddress, asked once per device with that device current (cached per host thread and device).
static int power_static_scratch(void** out) {
    static thread_local void* cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); set_error("power_iter_full_static: no current CUDA device"); return PDSC_ERR_DEVICE; }
    if (dev >= 0 && dev < 64 && cached[dev]) { *out = cached[dev]; return PDSC_OK; }
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_power_scratch);
    if (e != cudaSuccess || !p) {
        cudaGetLastError();
        set_error("power_iter_full_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
        return PDSC_ERR_CUDA;
    }
    if (dev >= 0 && dev < 64) cached[dev] = p;
    *out = p;
    return PDSC_OK;
}

int power_iter_full_static(const float* src, const float* tgt, float* v_out, int N, float sigma_spat, int num_iterations,
                           int32_t* iters_out, cudaStream_t st) {
    if (N > PDSC_POWER_ITER_STATIC_MAX_N) {
        set_error("power_iter_full_static: N=%d exceeds the static scratch (PDSC_POWER_ITER_STATIC_MAX_N=%d); use "
                  "pdsc_power_iter_full with a caller-supplied workspace", N, PDSC_POWER_ITER_STATIC_MAX_N);
        return PDSC_ERR_ARG;
    }
    void* scratch = nullptr;
    const int rc = power_static_scratch(&scratch);
    if (rc != PDSC_OK) return rc;
    return power_iter_full(src, tgt, v_out, N, sigma_spat, num_iterations, iters_out, scratch, sizeof(g_power_scratch), st);
}

}  // namespace pdsc
