// (N2) Putative-correspondence construction in descriptor space, the step right before PointDSC.forward.
// Replaces the numpy / torch CPU code of
//   datasets/LidarFeatureExtractor.py:109-118   distance = sqrt(2 - 2 S T^T + 1e-6), argmin per row, mutual check
//   algorithms/matching.py:22-65 (find_nn)      |a|^2 + |b|^2 - 2 a.b, clamp, sqrt, argmin (+ second nearest)
//   algorithms/matching.py:225-242 (nn_to_mutual) + torch_intersect :67-87
//
// nn_match_kernel: two query rows per thread (descriptors in registers), reference rows staged through shared
// memory in tiles and read as broadcasts; every (query, reference) dot product is ONE sequential fp32 FMA chain
// over the channels, so the A->B and the B->A pass see bit-identical values (the mutual check relies on that).
// Ties go to the lower index, like np.argmin / torch.min.  HBM traffic is the two descriptor sets (L2-resident
// at these sizes); the kernel is FP32-FMA bound: 2 Na Nb D flops.
// mutual_corr_kernel: keeps i with idx_ba[idx_ab[i]] == i (or every i) and compacts to [M,2] int64 in ascending i.
#include "common.cuh"

namespace pdsc {

constexpr int NM_THREADS = 128;
constexpr int NM_ROWS = 2;           // query rows per thread (NM_THREADS * NM_ROWS per CTA)
constexpr int NM_TILE = 64;          // reference rows per shared-memory tile

// The reference takes the argmin of sqrt(u) with
//   u = (2 - 2 dot) + 1e-6                        (np.sqrt(2 - 2 * dot + 1e-6): exact product, two rounded additions)
//   u = max((|a|^2 + |b|^2) - 2 dot, 1e-30)       (find_nn: clamp_min then sqrt)
// evaluated with exactly these rounding steps.  sqrt is monotone, so a candidate with u >= the current value can
// never win; only the rare candidate with a smaller u pays for the correctly rounded sqrt, which decides (two
// different u can round to the same distance, and then the lower index keeps the place, as in the reference).
__device__ __forceinline__ float nm_presqrt(float dot, float na, float nb, int metric) {
    if (metric == PDSC_MATCH_DOT) return __fadd_rn(__fsub_rn(2.0f, 2.0f * dot), 1e-6f);
    return fmaxf(__fsub_rn(__fadd_rn(na, nb), 2.0f * dot), 1e-30f);
}

template <int DPAD>
__global__ void __launch_bounds__(NM_THREADS) nn_match_kernel(const float* __restrict__ A, const float* __restrict__ Bm, int Na, int Nb,
                                                              int D, int metric, int64_t* __restrict__ idx1,
                                                              int64_t* __restrict__ idx2, float* __restrict__ val1,
                                                              float* __restrict__ val2) {
    __shared__ __align__(16) float tile[NM_TILE * DPAD];
    __shared__ float tnorm[NM_TILE];
    const int p = blockIdx.y;                           // problem (registration pair)
    A += (size_t)p * Na * D;
    Bm += (size_t)p * Nb * D;
    // NM_ROWS query rows per thread: every broadcast LDS.128 of a reference row feeds 4 * NM_ROWS FMAs (the
    // shared-memory pipe is per SM; at one row per thread it, not the FMA pipe, sets the pace)
    float a[NM_ROWS][DPAD];
    float na[NM_ROWS];
    int row[NM_ROWS];
#pragma unroll
    for (int q = 0; q < NM_ROWS; ++q) {
        row[q] = (blockIdx.x * NM_ROWS + q) * NM_THREADS + threadIdx.x;
        const int ic = min(row[q], Na - 1);
        na[q] = 0.f;
#pragma unroll
        for (int c = 0; c < DPAD; ++c) {
            a[q][c] = c < D ? A[(size_t)ic * D + c] : 0.f;
            na[q] = fmaf(a[q][c], a[q][c], na[q]);
        }
    }
    // the two smallest distances so far, their pre-sqrt values and indices
    float best[NM_ROWS], second[NM_ROWS], best_u[NM_ROWS], second_u[NM_ROWS];
    int ibest[NM_ROWS], isecond[NM_ROWS];
#pragma unroll
    for (int q = 0; q < NM_ROWS; ++q) { best[q] = second[q] = best_u[q] = second_u[q] = INFINITY; ibest[q] = isecond[q] = 0; }
    for (int j0 = 0; j0 < Nb; j0 += NM_TILE) {
        const int nj = min(NM_TILE, Nb - j0);
        __syncthreads();
        for (int t = threadIdx.x; t < NM_TILE * DPAD; t += NM_THREADS) {
            const int r = t / DPAD, c = t - r * DPAD;
            tile[t] = (r < nj && c < D) ? Bm[(size_t)(j0 + r) * D + c] : 0.f;
        }
        __syncthreads();
        if (threadIdx.x < NM_TILE) {
            float s = 0.f;
            for (int c = 0; c < DPAD; ++c) s = fmaf(tile[threadIdx.x * DPAD + c], tile[threadIdx.x * DPAD + c], s);
            tnorm[threadIdx.x] = s;
        }
        __syncthreads();
        for (int r = 0; r < nj; r += 4) {
            float d[NM_ROWS][4];
#pragma unroll
            for (int q = 0; q < NM_ROWS; ++q)
#pragma unroll
                for (int u = 0; u < 4; ++u) d[q][u] = 0.f;
#pragma unroll
            for (int c = 0; c < DPAD; c += 4) {
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const float4 b = *reinterpret_cast<const float4*>(&tile[(r + u) * DPAD + c]);   // rows past nj are zero
#pragma unroll
                    for (int q = 0; q < NM_ROWS; ++q) {
                        d[q][u] = fmaf(a[q][c], b.x, d[q][u]);
                        d[q][u] = fmaf(a[q][c + 1], b.y, d[q][u]);
                        d[q][u] = fmaf(a[q][c + 2], b.z, d[q][u]);
                        d[q][u] = fmaf(a[q][c + 3], b.w, d[q][u]);
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float nb = tnorm[r + u];
#pragma unroll
                for (int q = 0; q < NM_ROWS; ++q) {
                    const float uu = nm_presqrt(d[q][u], na[q], nb, metric);
                    if (uu < second_u[q] && r + u < nj) {
                        const float v = __fsqrt_rn(uu);
                        const int j = j0 + r + u;
                        if (v < best[q]) {
                            second[q] = best[q]; isecond[q] = ibest[q]; second_u[q] = best_u[q];
                            best[q] = v; ibest[q] = j; best_u[q] = uu;
                        } else if (v < second[q]) {
                            second[q] = v; isecond[q] = j; second_u[q] = uu;
                        }
                    }
                }
            }
        }
    }
#pragma unroll
    for (int q = 0; q < NM_ROWS; ++q) {
        if (row[q] < Na) {
            const size_t o = (size_t)p * Na + row[q];
            idx1[o] = ibest[q];
            if (idx2) idx2[o] = isecond[q];
            if (val1) val1[o] = best[q];
            if (val2) val2[o] = second[q];
        }
    }
}

// one CTA per problem; block-wide exclusive scan over chunks of 1024 rows keeps the output in ascending i
__global__ void __launch_bounds__(1024) mutual_corr_kernel(const int64_t* __restrict__ idx_ab, const int64_t* __restrict__ idx_ba,
                                                           int Na, int Nb, int use_mutual, int64_t* __restrict__ corr,
                                                           int32_t* __restrict__ count) {
    __shared__ int warp_tot[32];
    __shared__ int base;
    const int p = blockIdx.x;
    idx_ab += (size_t)p * Na;
    if (idx_ba) idx_ba += (size_t)p * Nb;
    corr += (size_t)p * Na * 2;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) base = 0;
    __syncthreads();
    for (int i0 = 0; i0 < Na; i0 += 1024) {
        const int i = i0 + threadIdx.x;
        int64_t j = 0;
        bool keep = false;
        if (i < Na) {
            j = idx_ab[i];
            keep = !use_mutual || (j >= 0 && j < Nb && idx_ba[j] == i);
        }
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        const int in_warp = __popc(m & ((1u << lane) - 1u));
        if (lane == 0) warp_tot[warp] = __popc(m);
        __syncthreads();
        int before = 0, total = 0;
        for (int w = 0; w < 32; ++w) {
            const int c = warp_tot[w];
            if (w < warp) before += c;
            total += c;
        }
        const int b = base;
        if (keep) {
            const int o = b + before + in_warp;
            corr[2 * o] = i;
            corr[2 * o + 1] = j;
        }
        __syncthreads();
        if (threadIdx.x == 0) base = b + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) count[p] = base;
}

int launch_nn_match(const float* A, const float* Bm, int P, int Na, int Nb, int D, int metric, int64_t* idx1, int64_t* idx2,
                    float* val1, float* val2, cudaStream_t st) {
    const dim3 grid((Na + NM_THREADS * NM_ROWS - 1) / (NM_THREADS * NM_ROWS), P);
    if (D <= 32) nn_match_kernel<32><<<grid, NM_THREADS, 0, st>>>(A, Bm, Na, Nb, D, metric, idx1, idx2, val1, val2);
    else if (D <= 36) nn_match_kernel<36><<<grid, NM_THREADS, 0, st>>>(A, Bm, Na, Nb, D, metric, idx1, idx2, val1, val2);
    else if (D <= 64) nn_match_kernel<64><<<grid, NM_THREADS, 0, st>>>(A, Bm, Na, Nb, D, metric, idx1, idx2, val1, val2);
    else { set_error("nn_match: descriptor width %d unsupported (at most 64)", D); return PDSC_ERR_ARG; }
    count_launches(1);
    return check_launch("nn_match");
}

int launch_mutual_corr(const int64_t* idx_ab, const int64_t* idx_ba, int P, int Na, int Nb, int use_mutual, int64_t* corr,
                       int32_t* count, cudaStream_t st) {
    mutual_corr_kernel<<<P, 1024, 0, st>>>(idx_ab, idx_ba, Na, Nb, use_mutual, corr, count);
    count_launches(1);
    return check_launch("mutual_corr");
}

}  // namespace pdsc
