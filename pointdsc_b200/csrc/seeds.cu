// (c) Seed selection: NMS local-maximum test with on-the-fly source distances + ordered top-S
// (PointDSC.pick_seeds, models/PointDSC.py:199-217), and feature-space kNN of the S seed rows
// only (knn, models/common.py:48-69 called at models/PointDSC.py:251-252).
// The reference builds three N x N boolean temporaries on top of the stored N x N src_dist and
// a full N x N Gram + topk over all N rows; here nothing of size N x N exists.
#include "common.cuh"

#include <math.h>
#include <stdlib.h>

namespace pdsc {

// ------------------------------------------------------------------------------------------
// pick_seeds, stage 1: suppressed[i] = exists j: !(score_i >= score_j) && !(|s_i-s_j| >= R)
// ------------------------------------------------------------------------------------------
constexpr int NMS_THREADS = 256;
constexpr int NMS_TILE = 1024;

__global__ void __launch_bounds__(NMS_THREADS) nms_kernel(const float* __restrict__ src,
                                                          const float* __restrict__ scores, int N, float T2,
                                                          int jchunk, int32_t* __restrict__ suppressed) {
    // two candidates per entry, coordinates negated: (-x_j, -x_j+1, -y_j, -y_j+1) and (-z_j, -z_j+1, score_j, score_j+1),
    // so that the distance runs on packed f32x2 instructions (each lane rounded like the scalar op: same bits)
    __shared__ float4 tile_a[NMS_TILE / 2], tile_b[NMS_TILE / 2];
    const int b = blockIdx.z;
    const float* sb = src + (size_t)b * N * 3;
    const float* cb = scores + (size_t)b * N;
    const int i = blockIdx.x * NMS_THREADS + threadIdx.x;
    float xi = 0.f, yi = 0.f, zi = 0.f, si = 0.f;
    if (i < N) { xi = sb[i * 3]; yi = sb[i * 3 + 1]; zi = sb[i * 3 + 2]; si = cb[i]; }
    const float2 xi2 = make_float2(xi, xi), yi2 = make_float2(yi, yi), zi2 = make_float2(zi, zi);
    const int j_begin = blockIdx.y * jchunk;
    const int j_end = min(N, j_begin + jchunk);
    bool keep = true;
    for (int j0 = j_begin; j0 < j_end; j0 += NMS_TILE) {
        const int nj = min(NMS_TILE, j_end - j0);
        const int nj2 = (nj + 1) >> 1;
        __syncthreads();
        for (int t = threadIdx.x; t < nj2; t += NMS_THREADS) {
            const int j = j0 + 2 * t;
            const bool two = 2 * t + 1 < nj;
            // an odd tail is padded with a point nothing is close to (d2 overflows to +inf) and the lowest score
            const float x1 = two ? sb[j * 3 + 3] : 3.0e38f, y1 = two ? sb[j * 3 + 4] : 3.0e38f, z1 = two ? sb[j * 3 + 5] : 3.0e38f;
            tile_a[t] = make_float4(-sb[j * 3], -x1, -sb[j * 3 + 1], -y1);
            tile_b[t] = make_float4(-sb[j * 3 + 2], -z1, cb[j], two ? cb[j + 1] : -INFINITY);
        }
        __syncthreads();
#pragma unroll 4
        for (int t = 0; t < nj2; ++t) {
            const float4 pa = tile_a[t], pb = tile_b[t];
            // sqrt(d2) >= R  <=>  d2 >= T2 (T2 = smallest fp32 whose correctly rounded sqrt reaches R, found on
            // the host), with d2 rounded exactly like torch.norm's mul/add sequence: no sqrt in the loop
            const float2 dx = __fadd2_rn(xi2, make_float2(pa.x, pa.y));
            const float2 dy = __fadd2_rn(yi2, make_float2(pa.z, pa.w));
            const float2 dz = __fadd2_rn(zi2, make_float2(pb.x, pb.y));
            // the sums stay scalar: ptxas (12.9) contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2, which would change the rounding
            const float2 qx = __fmul2_rn(dx, dx), qy = __fmul2_rn(dy, dy), qz = __fmul2_rn(dz, dz);
            const float2 d2 = make_float2(__fadd_rn(__fadd_rn(qx.x, qy.x), qz.x), __fadd_rn(__fadd_rn(qx.y, qy.y), qz.y));
            // suppressed by j  <=>  !(score_i >= score_j) && !(d2 >= T2)   (negated forms keep the reference's NaN behaviour)
            keep = keep && ((si >= pb.z) || (d2.x >= T2));
            keep = keep && ((si >= pb.w) || (d2.y >= T2));
        }
    }
    const bool sup = !keep;
    if (sup && i < N) suppressed[(size_t)b * N + i] = 1;
}

// ------------------------------------------------------------------------------------------
// pick_seeds, stage 1 on a uniform grid.  The test above is N^2 (190 us for 16 pairs x N=5000, FMA pipe 65 % busy: only
// fewer tests help).  A point j can only suppress i if |s_i - s_j| < R, so with cubic cells of edge c >= 1.001 R only the
// 27 cells around i can hold such a j; every candidate found there goes through EXACTLY the arithmetic of nms_kernel
// (same operations, same rounding, same negated comparisons), so the flags are bit-identical.  Points outside the 27 cells
// differ from i by more than c in some coordinate; the 0.1 % margin on c covers the fp32 rounding of the cell index and of
// d2 (both ~1e-7 relative).  Pairs with a non-finite coordinate keep the reference's "NaN suppresses everything below it"
// behaviour by scanning all j (flag set while the grid is built).
//   nms_grid_build_kernel  one CTA per pair: bounding box, cell edge c = max(1.001 R, extent / 32 * 1.001) (at most 32^3
//                          cells: the histogram lives in 128 KB of shared memory), counting sort of the points by cell
//                          (x fastest) -> records (x, y, z, score) in cell order + cell_start.
//   nms_grid_test_kernel   one thread per point: 3 x 3 runs of three x-adjacent cells (contiguous in the sorted order).
// ------------------------------------------------------------------------------------------
constexpr int NG_MAX = 32, NG_CELLS = NG_MAX * NG_MAX * NG_MAX, NG_THREADS = 1024;
struct NmsGridHeader { float minx, miny, minz, inv_c; int nx, ny, nz, brute; };

__device__ __forceinline__ int ng_cell(float v, float mn, float inv_c, int n) {
    const int c = (int)floorf((v - mn) * inv_c);
    return min(max(c, 0), n - 1);
}

__global__ void __launch_bounds__(NG_THREADS) nms_grid_build_kernel(const float* __restrict__ src, const float* __restrict__ scores,
                                                                    int N, float R, NmsGridHeader* __restrict__ hdr,
                                                                    int* __restrict__ cell_start, float4* __restrict__ rec) {
    extern __shared__ int hist[];                  // NG_CELLS counters, then (after the scan) running offsets
    __shared__ float red[6][32];
    __shared__ int warp_tot[32], bad;
    __shared__ NmsGridHeader h;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* sb = src + (size_t)b * N * 3;
    const float* cb = scores + (size_t)b * N;
    if (tid == 0) bad = 0;
    float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    bool nonfinite = false;
    for (int i = tid; i < N; i += NG_THREADS)
        for (int a = 0; a < 3; ++a) {
            const float v = sb[i * 3 + a];
            nonfinite |= !isfinite(v);
            mn[a] = fminf(mn[a], v); mx[a] = fmaxf(mx[a], v);
        }
    for (int a = 0; a < 3; ++a) {
        float lo = mn[a], hi = mx[a];
        for (int o = 16; o; o >>= 1) { lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o)); hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o)); }
        if (lane == 0) { red[a][warp] = lo; red[3 + a][warp] = hi; }
    }
    __syncthreads();
    if (nonfinite) bad = 1;
    for (int c = tid; c < NG_CELLS; c += NG_THREADS) hist[c] = 0;
    if (tid == 0) {
        float lo[3], hi[3];
        for (int a = 0; a < 3; ++a) {
            lo[a] = red[a][0]; hi[a] = red[3 + a][0];
            for (int w = 1; w < NG_THREADS / 32; ++w) { lo[a] = fminf(lo[a], red[a][w]); hi[a] = fmaxf(hi[a], red[3 + a][w]); }
        }
        const float ext = fmaxf(fmaxf(hi[0] - lo[0], hi[1] - lo[1]), hi[2] - lo[2]);
        float c = fmaxf(R * 1.001f, ext * (1.001f / NG_MAX));
        if (!(c > 0.f) || !isfinite(c)) c = 1.0f;
        h.minx = lo[0]; h.miny = lo[1]; h.minz = lo[2]; h.inv_c = 1.0f / c;
        h.nx = min(NG_MAX, (int)floorf((hi[0] - lo[0]) / c) + 1);
        h.ny = min(NG_MAX, (int)floorf((hi[1] - lo[1]) / c) + 1);
        h.nz = min(NG_MAX, (int)floorf((hi[2] - lo[2]) / c) + 1);
        h.brute = 0;
    }
    __syncthreads();
    if (tid == 0 && bad) { h.brute = 1; h.nx = h.ny = h.nz = 1; h.minx = h.miny = h.minz = 0.f; h.inv_c = 0.f; }
    __syncthreads();
    const NmsGridHeader g = h;
    const int ncell = g.nx * g.ny * g.nz;
    for (int i = tid; i < N; i += NG_THREADS) {
        const int c = (ng_cell(sb[i * 3 + 2], g.minz, g.inv_c, g.nz) * g.ny + ng_cell(sb[i * 3 + 1], g.miny, g.inv_c, g.ny)) * g.nx +
                      ng_cell(sb[i * 3], g.minx, g.inv_c, g.nx);
        atomicAdd(&hist[c], 1);
    }
    __syncthreads();
    // exclusive scan of hist[0, ncell): each thread owns a contiguous chunk
    const int per = (ncell + NG_THREADS - 1) / NG_THREADS, c0 = min(ncell, tid * per), c1 = min(ncell, c0 + per);
    int sum = 0;
    for (int c = c0; c < c1; ++c) sum += hist[c];
    int incl = sum;
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int t = warp_tot[lane], inc2 = t;
        for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, inc2, o); if (lane >= o) inc2 += u; }
        warp_tot[lane] = inc2 - t;
    }
    __syncthreads();
    int run = warp_tot[warp] + incl - sum;
    int* cs = cell_start + (size_t)b * (NG_CELLS + 1);
    for (int c = c0; c < c1; ++c) { const int n = hist[c]; hist[c] = run; cs[c] = run; run += n; }
    if (tid == 0) { cs[ncell] = N; hdr[b] = g; }
    __syncthreads();
    // scatter: the order inside a cell is whatever the atomics give -- the test is an OR over the candidates, so it does not matter
    float4* rb = rec + (size_t)b * N;
    for (int i = tid; i < N; i += NG_THREADS) {
        const float x = sb[i * 3], y = sb[i * 3 + 1], z = sb[i * 3 + 2];
        const int c = (ng_cell(z, g.minz, g.inv_c, g.nz) * g.ny + ng_cell(y, g.miny, g.inv_c, g.ny)) * g.nx + ng_cell(x, g.minx, g.inv_c, g.nx);
        rb[atomicAdd(&hist[c], 1)] = make_float4(x, y, z, cb[i]);
    }
}

__global__ void __launch_bounds__(256) nms_grid_test_kernel(const float* __restrict__ src, const float* __restrict__ scores, int N,
                                                            float T2, const NmsGridHeader* __restrict__ hdr,
                                                            const int* __restrict__ cell_start, const float4* __restrict__ rec,
                                                            int32_t* __restrict__ suppressed) {
    const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const NmsGridHeader g = hdr[b];
    const float xi = src[((size_t)b * N + i) * 3], yi = src[((size_t)b * N + i) * 3 + 1], zi = src[((size_t)b * N + i) * 3 + 2];
    const float si = scores[(size_t)b * N + i];
    const float4* rb = rec + (size_t)b * N;
    const int* cs = cell_start + (size_t)b * (NG_CELLS + 1);
    bool keep = true;
    auto test = [&](int j0, int j1) {
        for (int j = j0; j < j1; ++j) {
            const float4 q = rb[j];
            // exactly nms_kernel's arithmetic: d = x_i + (-x_j), squares, (qx + qy) + qz, all rounded to nearest
            const float dx = __fadd_rn(xi, -q.x), dy = __fadd_rn(yi, -q.y), dz = __fadd_rn(zi, -q.z);
            const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            keep = keep && ((si >= q.w) || (d2 >= T2));
        }
    };
    if (g.brute) {
        test(0, N);
    } else {
        const int cx = ng_cell(xi, g.minx, g.inv_c, g.nx), cy = ng_cell(yi, g.miny, g.inv_c, g.ny), cz = ng_cell(zi, g.minz, g.inv_c, g.nz);
        const int x0 = max(cx - 1, 0), x1 = min(cx + 1, g.nx - 1);
        for (int z = max(cz - 1, 0); z <= min(cz + 1, g.nz - 1); ++z)
            for (int y = max(cy - 1, 0); y <= min(cy + 1, g.ny - 1); ++y) {
                const int base = (z * g.ny + y) * g.nx;
                test(cs[base + x0], cs[base + x1 + 1]);
            }
    }
    if (!keep) suppressed[(size_t)b * N + i] = 1;
}

// ------------------------------------------------------------------------------------------
// pick_seeds, stage 2: order by (key descending, index ascending), key = score*is_local_max.
// One CTA per pair, bitonic sort of 64-bit composite keys (all distinct -> deterministic).
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void bitonic_sort(unsigned long long* a, int n_pow2) {
    for (int k = 2; k <= n_pow2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            __syncthreads();
            for (int t = threadIdx.x; t < n_pow2; t += blockDim.x) {
                const int p = t ^ j;
                if (p > t) {
                    const unsigned long long x = a[t], y = a[p];
                    const bool up = (t & k) == 0;
                    if ((x > y) == up) { a[t] = y; a[p] = x; }
                }
            }
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(1024) seed_order_kernel(const float* __restrict__ scores,
                                                          const int32_t* __restrict__ suppressed, int N, int S,
                                                          int n_pow2, unsigned long long* __restrict__ gbuf,
                                                          int64_t* __restrict__ seeds,
                                                          uint8_t* __restrict__ is_local_max,
                                                          int32_t* __restrict__ n_pos) {
    extern __shared__ unsigned long long sbuf[];
    __shared__ unsigned long long cand[1024];     // thread minima, then the candidates of the short sort
    __shared__ int pos_count, n_cand;
    const int b = blockIdx.x;
    unsigned long long* a = gbuf ? gbuf + (size_t)b * n_pow2 : sbuf;
    if (threadIdx.x == 0) { pos_count = 0; n_cand = 0; }
    __syncthreads();
    int local_pos = 0;
    unsigned long long my_min = ~0ull;
    for (int i = threadIdx.x; i < n_pow2; i += blockDim.x) {
        unsigned long long key = ~0ull;  // padding sorts last
        if (i < N) {
            const bool lm = suppressed[(size_t)b * N + i] == 0;
            float kf = scores[(size_t)b * N + i] * (lm ? 1.0f : 0.0f);
            if (kf == 0.0f) kf = 0.0f;  // -0 and +0 are equal keys in the reference
            local_pos += kf > 0.0f ? 1 : 0;
            key = ((unsigned long long)(~float_to_ordered(kf)) << 32) | (unsigned long long)(uint32_t)i;
            if (is_local_max) is_local_max[(size_t)b * N + i] = lm ? 1 : 0;
        }
        a[i] = key;
        my_min = key < my_min ? key : my_min;
    }
    local_pos = warp_sum(local_pos);
    if ((threadIdx.x & 31) == 0 && local_pos) atomicAdd(&pos_count, local_pos);
    // Short path: only the S first keys are wanted.  The S-th smallest of the 1024 per-thread minima is an upper bound of
    // the S-th smallest key (the minima are S distinct keys at or below it); the few keys at or below the bound
    // (about S * 1.3) are sorted instead of all n_pow2.  Same result as the full sort: keys are distinct.
    bool done = false;
    if (blockDim.x == 1024 && S <= 1024 && S <= N) {
        cand[threadIdx.x] = my_min;
        bitonic_sort(cand, 1024);
        const unsigned long long bound = cand[S - 1];
        __syncthreads();                                   // everyone has the bound: cand may be overwritten
        if (bound != ~0ull) {
            for (int i = threadIdx.x; i < n_pow2; i += blockDim.x) {
                const unsigned long long key = a[i];
                if (key <= bound) {
                    const int pos = atomicAdd(&n_cand, 1);
                    if (pos < 1024) cand[pos] = key;
                }
            }
            __syncthreads();
            const int nc = n_cand;
            if (nc <= 1024) {
                if ((int)threadIdx.x >= nc) cand[threadIdx.x] = ~0ull;
                bitonic_sort(cand, 1024);
                for (int s = threadIdx.x; s < S; s += blockDim.x) seeds[(size_t)b * S + s] = (int64_t)(cand[s] & 0xffffffffull);
                done = true;
            }
        }
    }
    if (!done) {
        bitonic_sort(a, n_pow2);
        for (int s = threadIdx.x; s < S; s += blockDim.x) seeds[(size_t)b * S + s] = (int64_t)(a[s] & 0xffffffffull);
    }
    if (threadIdx.x == 0 && n_pos) n_pos[b] = pos_count;
}

static int next_pow2(int n) {
    int p = 1;
    while (p < n) p <<= 1;
    return p;
}

constexpr int SORT_SMEM_MAX_ELEMS = 16384;  // 128 KB of 64-bit keys

// workspace: suppressed flags | (large N) global sort buffer | grid: headers, cell_start, point records
static size_t ps_sort_off(int B, int N) { return align_up(sizeof(int32_t) * (size_t)B * N, 256); }
static size_t ps_grid_off(int B, int N) {
    const int p2 = next_pow2(N);
    return ps_sort_off(B, N) + (p2 > SORT_SMEM_MAX_ELEMS ? align_up(sizeof(unsigned long long) * (size_t)B * p2, 256) : 0);
}
size_t pick_seeds_workspace(int B, int N) {
    return ps_grid_off(B, N) + align_up(sizeof(NmsGridHeader) * (size_t)B, 256) + align_up(sizeof(int) * (size_t)B * (NG_CELLS + 1), 256) +
           sizeof(float4) * (size_t)B * N;
}

int launch_pick_seeds(const float* src, const float* scores, int B, int N, float R, int S, int64_t* seeds,
                      uint8_t* is_local_max, int32_t* n_pos, void* ws, size_t ws_bytes, cudaStream_t st) {
    if (ws_bytes < pick_seeds_workspace(B, N)) {
        set_error("pick_seeds: workspace too small (%zu < %zu)", ws_bytes, pick_seeds_workspace(B, N));
        return PDSC_ERR_WORKSPACE;
    }
    int32_t* suppressed = reinterpret_cast<int32_t*>(ws);
    cudaMemsetAsync(suppressed, 0, sizeof(int32_t) * (size_t)B * N, st);
    const int iblocks = (N + NMS_THREADS - 1) / NMS_THREADS;
    // enough j-splits to fill the machine a few times over
    int jsplit = (4 * num_sms() + iblocks * B - 1) / (iblocks * B);
    jsplit = max(1, min(jsplit, (N + NMS_TILE - 1) / NMS_TILE));
    int jchunk = (N + jsplit - 1) / jsplit;
    jchunk = (jchunk + NMS_TILE - 1) / NMS_TILE * NMS_TILE;
    jsplit = (N + jchunk - 1) / jchunk;
    // squared-distance threshold equivalent to the reference's `norm >= R` (sqrtf is monotone and correctly rounded)
    const float T2 = squared_threshold(R);
    // PDSC_NMS_BRUTE=1 (tuning build): the N^2 kernel, which the tests compare the grid kernels with, bit for bit
    if (tuning_int("PDSC_NMS_BRUTE", 0) != 0 || !(R > 0.f)) {
        nms_kernel<<<dim3(iblocks, jsplit, B), NMS_THREADS, 0, st>>>(src, scores, N, T2, jchunk, suppressed);
    } else {
        char* gbase = reinterpret_cast<char*>(ws) + ps_grid_off(B, N);
        NmsGridHeader* hdr = reinterpret_cast<NmsGridHeader*>(gbase);
        int* cell_start = reinterpret_cast<int*>(gbase + align_up(sizeof(NmsGridHeader) * (size_t)B, 256));
        float4* rec = reinterpret_cast<float4*>(reinterpret_cast<char*>(cell_start) + align_up(sizeof(int) * (size_t)B * (NG_CELLS + 1), 256));
        opt_in_smem<nms_grid_build_kernel>((int)(sizeof(int) * NG_CELLS));
        nms_grid_build_kernel<<<B, NG_THREADS, sizeof(int) * NG_CELLS, st>>>(src, scores, N, R, hdr, cell_start, rec);
        nms_grid_test_kernel<<<dim3((N + 255) / 256, B), 256, 0, st>>>(src, scores, N, T2, hdr, cell_start, rec, suppressed);
        count_launches(1);
    }
    const int p2 = next_pow2(N);
    unsigned long long* gbuf = nullptr;
    size_t smem = sizeof(unsigned long long) * (size_t)p2;
    if (p2 > SORT_SMEM_MAX_ELEMS) {
        gbuf = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(ws) + ps_sort_off(B, N));
        smem = 0;
    } else {
        opt_in_smem<seed_order_kernel>((int)(sizeof(unsigned long long) * SORT_SMEM_MAX_ELEMS));
    }
    seed_order_kernel<<<B, 1024, smem, st>>>(scores, suppressed, N, S, p2, gbuf, seeds, is_local_max, n_pos);
    count_launches(2);
    return check_launch("pick_seeds");
}

// ------------------------------------------------------------------------------------------
// seed kNN, stage 1: D[b][s][j] = 2 - 2 <f_seed(s), f_j>  (self -> +inf), fp32 FMA GEMM
// CTA tile 64 seeds x 256 candidates, 8 x 8 outputs per thread, channels in chunks of 32.  Both operands stay
// row-major in shared memory ([row][channel], pitch 36): a thread reads 4 consecutive channels of a row with one
// LDS.128 (seed rows: warp-wide broadcast; candidate rows tx + 32 j: conflict free), i.e. 4 LDS.128 per 64 FMAs,
// and every dot product is one sequential FMA chain over the channels.
// ------------------------------------------------------------------------------------------
constexpr int KG_TS = 64, KG_TJ = 256, KG_TC = 32, KG_THREADS = 256, KG_PITCH = KG_TC + 4;

__global__ void __launch_bounds__(KG_THREADS, 2) knn_dist_kernel(const float* __restrict__ feat,
                                                                 const int64_t* __restrict__ seeds, int N, int C, int S,
                                                                 float* __restrict__ D) {
    __shared__ __align__(16) float As[KG_TS * KG_PITCH];
    __shared__ __align__(16) float Bs[KG_TJ * KG_PITCH];
    __shared__ int seed_idx[KG_TS];
    const int b = blockIdx.z;
    const int s0 = blockIdx.y * KG_TS, j0 = blockIdx.x * KG_TJ;
    const float* fb = feat + (size_t)b * N * C;
    if (threadIdx.x < KG_TS) {
        const int s = s0 + threadIdx.x;
        seed_idx[threadIdx.x] = s < S ? (int)seeds[(size_t)b * S + s] : -1;
    }
    __syncthreads();
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // seeds ty*8 + i, candidates tx + 32 j
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    for (int c0 = 0; c0 < C; c0 += KG_TC) {
#pragma unroll
        for (int rep = 0; rep < KG_TS * (KG_TC / 4) / KG_THREADS; ++rep) {   // A: 64 seeds x 8 float4
            const int f = threadIdx.x + rep * KG_THREADS;
            const int r = f >> 3, c4 = f & 7;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            const int si = seed_idx[r];
            if (si >= 0) v = *reinterpret_cast<const float4*>(fb + (size_t)si * C + c0 + c4 * 4);
            *reinterpret_cast<float4*>(&As[r * KG_PITCH + c4 * 4]) = v;
        }
#pragma unroll
        for (int rep = 0; rep < KG_TJ * (KG_TC / 4) / KG_THREADS; ++rep) {   // B: 256 candidates x 8 float4
            const int f = threadIdx.x + rep * KG_THREADS;
            const int r = f >> 3, c4 = f & 7;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (j0 + r < N) v = *reinterpret_cast<const float4*>(fb + (size_t)(j0 + r) * C + c0 + c4 * 4);
            *reinterpret_cast<float4*>(&Bs[r * KG_PITCH + c4 * 4]) = v;
        }
        __syncthreads();
#pragma unroll 1
        for (int c = 0; c < KG_TC; c += 4) {
            float4 a[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = *reinterpret_cast<const float4*>(&As[(ty * 8 + i) * KG_PITCH + c]);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 bq = *reinterpret_cast<const float4*>(&Bs[(tx + 32 * j) * KG_PITCH + c]);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    acc[i][j] = fmaf(a[i].x, bq.x, acc[i][j]);
                    acc[i][j] = fmaf(a[i].y, bq.y, acc[i][j]);
                    acc[i][j] = fmaf(a[i].z, bq.z, acc[i][j]);
                    acc[i][j] = fmaf(a[i].w, bq.w, acc[i][j]);
                }
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int s = s0 + ty * 8 + i;
        if (s >= S) continue;
        const int self = seed_idx[ty * 8 + i];
        float* row = D + ((size_t)b * S + s) * N;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int jj = j0 + tx + 32 * j;
            if (jj < N) row[jj] = (jj == self) ? INFINITY : __fsub_rn(2.0f, __fmul_rn(2.0f, acc[i][j]));   // common.py:58-60
        }
    }
}

// ------------------------------------------------------------------------------------------
// seed kNN, stage 2: exact k smallest (dist, index) per seed row.  One CTA per row:
//   * the row is read from global memory ONCE, as ordered keys, into shared memory (rows of up to KS_SMEM_KEYS
//     elements; longer rows are re-read from global memory, which L1/L2 serve); every thread keeps the minimum of
//     the elements it staged;
//   * fast path: an upper bound of the k-th smallest element from the thread minima (every warp takes the
//     ceil(k/4)-th smallest of its 32 minima, the largest of the four has at least k distinct elements at or below
//     it); only ~k * 1.4 elements of a row lie at or below the bound: they are compacted and rank-sorted by
//     (key, index) directly -- exact, no histogram;
//   * slow path (more than KS_CAND elements at or below the bound: heavy ties): 4-pass radix select on the keys.  Distances 2 - 2 cos sit in a narrow band, so the first passes put nearly every
//     element into one bin: votes are aggregated per warp (__match_any_sync) before the shared-memory atomic;
//   * one ordered pass: every thread owns a contiguous segment of the row, a block scan of the per-thread counts
//     gives each thread its output slots ("< T" first, then the "== T" elements in index order);
//   * rank sort of the k winners by (key, index).
// ------------------------------------------------------------------------------------------
constexpr int KS_THREADS = 128;
constexpr int KS_MAXK = 64;
constexpr int KS_SMEM_KEYS = 16384;
constexpr int KS_CAND = 256;

template <bool SMEM>
__global__ void __launch_bounds__(KS_THREADS) knn_select_kernel(const float* __restrict__ D, int N, int S, int k,
                                                                int64_t* __restrict__ knn_idx,
                                                                float* __restrict__ knn_dist) {
    extern __shared__ unsigned int keys[];          // SMEM: the row as ordered keys
    __shared__ unsigned int hist[256];
    __shared__ unsigned int sh_prefix, sh_want;
    __shared__ unsigned int cand_key[KS_CAND];
    __shared__ int cand_idx[KS_CAND];
    __shared__ unsigned int warp_bound[KS_THREADS / 32];
    __shared__ unsigned int sh_bound;
    __shared__ int n_cand;
    __shared__ int warp_less[KS_THREADS / 32], warp_eq[KS_THREADS / 32];
    const int row_id = blockIdx.x;   // b*S + s
    const float* row = D + (size_t)row_id * N;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    auto key_at = [&](int j) -> unsigned { return SMEM ? keys[j] : float_to_ordered(__ldg(row + j)); };

    unsigned my_min = 0xffffffffu;                   // above the key of +inf: a thread without elements ranks last
#pragma unroll 8
    for (int j = tid; j < N; j += KS_THREADS) {
        const unsigned key = float_to_ordered(__ldcs(row + j));
        if (SMEM) keys[j] = key;
        my_min = min(my_min, key);
    }
    if (tid == 0) { sh_prefix = 0u; sh_want = (unsigned)k; n_cand = 0; }
    {
        // every warp finds the q-th smallest of its 32 minima (rank by counting over shuffles, ties broken by lane),
        // q = ceil(k / warps): the largest of the four has at least 4 q >= k elements at or below it
        int rank = 0;
#pragma unroll 8
        for (int o = 0; o < 32; ++o) {
            const unsigned m = __shfl_sync(0xffffffffu, my_min, o);
            rank += (m < my_min || (m == my_min && o < lane)) ? 1 : 0;
        }
        constexpr int NW = KS_THREADS / 32;
        if (rank == (k + NW - 1) / NW - 1) warp_bound[warp] = my_min;
    }
    __syncthreads();
    if (tid == 0) {
        unsigned bmax = warp_bound[0];
        for (int w = 1; w < KS_THREADS / 32; ++w) bmax = max(bmax, warp_bound[w]);
        sh_bound = bmax;
    }
    __syncthreads();
    {
        const unsigned bound = sh_bound;
        if (my_min <= bound) {                       // only threads whose minimum passes can hold candidates
            for (int j = tid; j < N; j += KS_THREADS) {
                const unsigned key = key_at(j);
                if (key <= bound) {
                    const int p = atomicAdd(&n_cand, 1);
                    if (p < KS_CAND) { cand_key[p] = key; cand_idx[p] = j; }
                }
            }
        }
    }
    __syncthreads();
    const int nc = n_cand;
    if (nc <= KS_CAND) {
        // rank sort of the candidates by (key, index); the first k ranks are the answer
        for (int c = tid; c < nc; c += KS_THREADS) {
            const unsigned mk = cand_key[c];
            const int mi = cand_idx[c];
            int rank = 0;
            for (int o = 0; o < nc; ++o) {
                const unsigned ok = cand_key[o];
                const int oi = cand_idx[o];
                rank += (ok < mk || (ok == mk && oi < mi)) ? 1 : 0;
            }
            if (rank < k) {
                knn_idx[(size_t)row_id * k + rank] = mi;
                if (knn_dist) knn_dist[(size_t)row_id * k + rank] = row[mi];
            }
        }
        return;
    }
    // ---- slow path: radix select ----
    const int n_round = (N + KS_THREADS - 1) / KS_THREADS * KS_THREADS;     // whole warps take part in every vote
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 24 - 8 * pass;
        for (int i = tid; i < 256; i += KS_THREADS) hist[i] = 0u;
        __syncthreads();
        const unsigned prefix = sh_prefix;
        const unsigned himask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
        for (int j = tid; j < n_round; j += KS_THREADS) {
            unsigned bin = 0xffffu;                                         // "no vote"
            if (j < N) {
                const unsigned key = key_at(j);
                if ((key & himask) == prefix) bin = (key >> shift) & 255u;
            }
            const unsigned peers = __match_any_sync(0xffffffffu, bin);
            if (bin != 0xffffu && lane == __ffs(peers) - 1) atomicAdd(&hist[bin], (unsigned)__popc(peers));
        }
        __syncthreads();
        if (warp == 0) {
            // bin of the want-th element: inclusive prefix sums of the 256 counts, 8 bins per lane
            unsigned c[8], tot = 0u;
#pragma unroll
            for (int i = 0; i < 8; ++i) { c[i] = hist[lane * 8 + i]; tot += c[i]; }
            unsigned incl = tot;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned o = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += o;
            }
            const unsigned want = sh_want;
            unsigned cum = incl - tot;                                      // elements in the bins of lower lanes
            const bool here = cum < want && want <= incl;                   // exactly one lane
            if (here) {
                int i = 0;
                for (; i < 7; ++i) {
                    if (cum + c[i] >= want) break;
                    cum += c[i];
                }
                sh_want = want - cum;
                sh_prefix = prefix | ((unsigned)(lane * 8 + i) << shift);
            }
        }
        __syncthreads();
    }
    const unsigned T = sh_prefix;           // key of the k-th smallest element
    const int want_eq = (int)sh_want;       // how many elements equal to T belong to the answer
    const int base_eq = k - want_eq;        // == number of elements with key < T
    // ordered pass over contiguous segments (odd length: conflict-free shared-memory reads)
    const int seg = ((N + KS_THREADS - 1) / KS_THREADS) | 1;
    const int j_lo = min(tid * seg, N), j_hi = min(j_lo + seg, N);
    int n_less = 0, n_eq = 0;
    for (int j = j_lo; j < j_hi; ++j) {
        const unsigned key = key_at(j);
        n_less += key < T ? 1 : 0;
        n_eq += key == T ? 1 : 0;
    }
    int less_incl = n_less, eq_incl = n_eq;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int a = __shfl_up_sync(0xffffffffu, less_incl, d), b = __shfl_up_sync(0xffffffffu, eq_incl, d);
        if (lane >= d) { less_incl += a; eq_incl += b; }
    }
    if (lane == 31) { warp_less[warp] = less_incl; warp_eq[warp] = eq_incl; }
    __syncthreads();
    int less_off = less_incl - n_less, eq_off = eq_incl - n_eq;
    for (int w = 0; w < warp; ++w) { less_off += warp_less[w]; eq_off += warp_eq[w]; }
    if (n_less > 0 || (n_eq > 0 && eq_off < want_eq)) {
        for (int j = j_lo; j < j_hi; ++j) {
            const unsigned key = key_at(j);
            if (key < T) {
                cand_key[less_off] = key;
                cand_idx[less_off] = j;
                ++less_off;
            } else if (key == T) {
                if (eq_off < want_eq) {
                    cand_key[base_eq + eq_off] = key;
                    cand_idx[base_eq + eq_off] = j;
                }
                ++eq_off;
            }
        }
    }
    __syncthreads();
    // rank sort by (key, index)
    if (tid < k) {
        const unsigned mk = cand_key[tid];
        const int mi = cand_idx[tid];
        int rank = 0;
        for (int o = 0; o < k; ++o) {
            const unsigned ok = cand_key[o];
            const int oi = cand_idx[o];
            rank += (ok < mk || (ok == mk && oi < mi)) ? 1 : 0;
        }
        knn_idx[(size_t)row_id * k + rank] = mi;
        if (knn_dist) knn_dist[(size_t)row_id * k + rank] = row[mi];
    }
}

size_t knn_tc_image_bytes(int B, int S);                                                                        // knn_tc.cu
int launch_knn_gram_tc(const float* normed, const int64_t* seeds, int B, int N, int S, float* D, void* img, cudaStream_t st);

// distances [B][S][N] | seed operand images of the tensor-core Gram
size_t seed_knn_workspace(int B, int N, int S) { return align_up(sizeof(float) * (size_t)B * S * N, 1024) + knn_tc_image_bytes(B, S); }

int launch_seed_knn(const float* normed, const int64_t* seeds, int B, int N, int C, int S, int k,
                    int64_t* knn_idx, float* knn_dist, void* ws, size_t ws_bytes, cudaStream_t st) {
    if (k > KS_MAXK || k < 1 || k > N - 1) {
        set_error("seed_knn: k=%d unsupported (1..%d, < N)", k, KS_MAXK);
        return PDSC_ERR_ARG;
    }
    if (C % KG_TC != 0) {
        set_error("seed_knn: C=%d must be a multiple of %d", C, KG_TC);
        return PDSC_ERR_ARG;
    }
    if (ws_bytes < seed_knn_workspace(B, N, S)) {
        set_error("seed_knn: workspace too small (%zu < %zu)", ws_bytes, seed_knn_workspace(B, N, S));
        return PDSC_ERR_WORKSPACE;
    }
    float* D = reinterpret_cast<float*>(ws);
    // C == 128: Gram on the tensor cores (knn_tc.cu); PDSC_KNN_FP32=1 (tuning build) keeps the fp32 CUDA-core Gram (tests
    // compare the two)
    const bool force_fp32 = tuning_int("PDSC_KNN_FP32", 0) != 0;
    if (C == 128 && !force_fp32) {
        void* img = reinterpret_cast<char*>(ws) + align_up(sizeof(float) * (size_t)B * S * N, 1024);
        const int rc = launch_knn_gram_tc(normed, seeds, B, N, S, D, img, st);
        if (rc != PDSC_OK) return rc;
    } else {
        dim3 grid((N + KG_TJ - 1) / KG_TJ, (S + KG_TS - 1) / KG_TS, B);
        knn_dist_kernel<<<grid, KG_THREADS, 0, st>>>(normed, seeds, N, C, S, D);
        count_launches(1);
    }
    if (N <= KS_SMEM_KEYS) {
        const int smem = (int)sizeof(unsigned int) * N;
        if (smem > 40 * 1024) opt_in_smem<knn_select_kernel<true>>((int)sizeof(unsigned int) * KS_SMEM_KEYS);
        knn_select_kernel<true><<<B * S, KS_THREADS, smem, st>>>(D, N, S, k, knn_idx, knn_dist);
    } else {
        knn_select_kernel<false><<<B * S, KS_THREADS, 0, st>>>(D, N, S, k, knn_idx, knn_dist);
    }
    count_launches(1);
    return check_launch("seed_knn");
}

}  // namespace pdsc
