// (d) Per-seed local spectral matching and weighted Kabsch.
// Replaces cal_seed_trans models/PointDSC.py:257-320 (gathers, [S,k,k] feature and spatial
// consistency matrices), cal_leading_eigenvector :338-358 (power iteration with a host sync per
// step) and rigid_transform_3d models/common.py:7-45 (diag_embed, D2H, LAPACK 3x3 SVD, H2D).
// One CTA per seed keeps the k x C feature block, the k x k matrix and all iterates in shared
// memory; nothing but the 4x4 transform leaves the SM.
//
// Stopping rule: the reference tests torch.allclose over the WHOLE [S,k,1] tensor, so all seeds
// of a pair stop together.  Every seed runs all T iterations, records in a per-pair bit mask
// at which iterations it had NOT converged, and speculatively finishes with the last iterate
// (the common case: some seed never converges within T).  A second, normally idle kernel redoes
// the Kabsch step from the stored iterate when the pair did stop early.
#include "common.cuh"

namespace pdsc {

constexpr int HY_THREADS = 128;
constexpr int HY_MAXK = 64;

struct HypoSmem {
    int idx[HY_MAXK];
    float src[HY_MAXK][3];
    float tgt[HY_MAXK][3];
};

// Weighted Kabsch for one seed by one warp.  v: un-normalised eigenvector (smem or global).
__device__ void seed_kabsch_warp(const float* v, const float (*ps)[3], const float (*pt)[3], int k, float* T_out,
                                 float* w_out) {
    const int lane = threadIdx.x & 31;
    float vv[2], a[2][3], bq[2][3];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int r = lane + 32 * h;
        const bool ok = r < k;
        vv[h] = ok ? v[r] : 0.f;
#pragma unroll
        for (int d = 0; d < 3; ++d) { a[h][d] = ok ? ps[r][d] : 0.f; bq[h][d] = ok ? pt[r][d] : 0.f; }
    }
    // total_weight / (sum + 1e-6)   (models/PointDSC.py:282)
    const float vsum = (float)warp_sum((double)vv[0] + (double)vv[1]);
    float w[2];
    w[0] = __fdiv_rn(vv[0], vsum + 1e-6f);
    w[1] = __fdiv_rn(vv[1], vsum + 1e-6f);
    if (w_out) {
        if (lane < k) w_out[lane] = w[0];
        if (lane + 32 < k) w_out[lane + 32] = w[1];
    }
    // rigid_transform_3d: weights < 0 -> 0 (common.py:20), centroids (:24-25)
    w[0] = w[0] < 0.f ? 0.f : w[0];
    w[1] = w[1] < 0.f ? 0.f : w[1];
    double acc[7];
    acc[0] = (double)w[0] + (double)w[1];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        acc[1 + d] = (double)__fmul_rn(a[0][d], w[0]) + (double)__fmul_rn(a[1][d], w[1]);
        acc[4 + d] = (double)__fmul_rn(bq[0][d], w[0]) + (double)__fmul_rn(bq[1][d], w[1]);
    }
#pragma unroll
    for (int i = 0; i < 7; ++i) acc[i] = warp_sum(acc[i]);
    const float ws = (float)acc[0] + 1e-6f;
    float cA[3], cB[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) { cA[d] = __fdiv_rn((float)acc[1 + d], ws); cB[d] = __fdiv_rn((float)acc[4 + d], ws); }
    double H[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            double h = 0.0;
#pragma unroll
            for (int hh = 0; hh < 2; ++hh)
                h += (double)w[hh] * (double)(a[hh][i] - cA[i]) * (double)(bq[hh][j] - cB[j]);
            H[i * 3 + j] = warp_sum(h);
        }
    if (lane == 0) {
        double R[9];
        kabsch_rotation(H, R);
        write_transform(T_out, R, cA, cB);
    }
}

__global__ void __launch_bounds__(HY_THREADS) seed_hypo_kernel(
    const float* __restrict__ feat, const float* __restrict__ src, const float* __restrict__ tgt,
    const int64_t* __restrict__ knn_idx, int N, int C, int S, int k, float sig2_feat, float sig2_spat, int T,
    float* __restrict__ seed_trans, float* __restrict__ weights, float* __restrict__ iterates,
    unsigned int* __restrict__ notconv_mask) {
    extern __shared__ __align__(16) float dyn[];
    __shared__ HypoSmem sm;
    const int kp = (k + 3) & ~3;        // rows padded to a multiple of 4
    const int ldf = C + 4;              // feature row stride (floats), keeps float4 alignment
    const int ldm = kp + 1;             // odd row stride: conflict-free column walks
    float* F = dyn;                     // [kp][ldf]
    float* M = F + (size_t)kp * ldf;    // [k][ldm]
    float* V = M + (size_t)kp * ldm;    // [T+1][HY_MAXK]
    const int row_id = blockIdx.x;      // b*S + s
    const int b = row_id / S;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* fb = feat + (size_t)b * N * C;

    if (tid < k) {
        const int j = (int)knn_idx[(size_t)row_id * k + tid];
        sm.idx[tid] = j;
        const float* s = src + ((size_t)b * N + j) * 3;
        const float* t = tgt + ((size_t)b * N + j) * 3;
        sm.src[tid][0] = s[0]; sm.src[tid][1] = s[1]; sm.src[tid][2] = s[2];
        sm.tgt[tid][0] = t[0]; sm.tgt[tid][1] = t[1]; sm.tgt[tid][2] = t[2];
    }
    __syncthreads();
    // gather the k feature rows (knn_features, models/PointDSC.py:257); one warp per row
    const int c4n = C >> 2;
    for (int r = warp; r < kp; r += HY_THREADS / 32) {
        float4* dst = reinterpret_cast<float4*>(F + (size_t)r * ldf);
        if (r < k) {
            const float4* srow = reinterpret_cast<const float4*>(fb + (size_t)sm.idx[r] * C);
            for (int c = lane; c < c4n; c += 32) dst[c] = srow[c];
        } else {
            for (int c = lane; c < c4n; c += 32) dst[c] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    __syncthreads();
    // total_knn_M = feature_knn_M * spatial_knn_M, zero diagonal (models/PointDSC.py:258-278).
    // The k x k matrix is cut into upper-triangular 4 x 4 blocks, mirrored on store; a block is two work items of
    // 2 x 4 outputs, so that all four warps take part.  Block q owns rows q, q + kb, q + 2 kb, q + 3 kb: neighbouring
    // threads (neighbouring bj) then read neighbouring feature rows, which the row pitch of C + 4 floats spreads over
    // all banks (rows 4 q .. 4 q + 3 per block would put every second block on the same banks).
    const int kb = kp >> 2;
    const int nblk = kb * (kb + 1) / 2;
    for (int item = tid; item < 2 * nblk; item += HY_THREADS) {
        const int blk = item >> 1, half = item & 1;
        int bi = 0, rem = blk;
        while (rem >= kb - bi) { rem -= kb - bi; ++bi; }
        const int bj = bi + rem;
        float acc[2][4] = {};
        const float* Fa = F + (size_t)(bi + kb * 2 * half) * ldf;
        const float* Fb = F + (size_t)bj * ldf;
        const size_t rstep = (size_t)kb * ldf;
        for (int c = 0; c < C; c += 4) {
            float4 av[2], bv[4];
#pragma unroll
            for (int i = 0; i < 2; ++i) av[i] = *reinterpret_cast<const float4*>(Fa + i * rstep + c);
#pragma unroll
            for (int j = 0; j < 4; ++j) bv[j] = *reinterpret_cast<const float4*>(Fb + j * rstep + c);
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    acc[i][j] = fmaf(av[i].x, bv[j].x, acc[i][j]);
                    acc[i][j] = fmaf(av[i].y, bv[j].y, acc[i][j]);
                    acc[i][j] = fmaf(av[i].z, bv[j].z, acc[i][j]);
                    acc[i][j] = fmaf(av[i].w, bv[j].w, acc[i][j]);
                }
        }
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int ra = bi + kb * (2 * half + i), rb = bj + kb * j;
                if (ra >= k || rb >= k) continue;
                float m = 0.f;
                if (ra != rb) {
                    const float fm = fmaxf(__fsub_rn(1.0f, __fdiv_rn(__fsub_rn(1.0f, acc[i][j]), sig2_feat)), 0.0f);
                    const float ds = dist3(sm.src[ra][0], sm.src[ra][1], sm.src[ra][2], sm.src[rb][0], sm.src[rb][1], sm.src[rb][2]);
                    const float dt = dist3(sm.tgt[ra][0], sm.tgt[ra][1], sm.tgt[ra][2], sm.tgt[rb][0], sm.tgt[rb][1], sm.tgt[rb][2]);
                    m = __fmul_rn(fm, compat_exact(ds, dt, sig2_spat));
                }
                M[ra * ldm + rb] = m;
                M[rb * ldm + ra] = m;
            }
    }
    __syncthreads();
    if (warp != 0) return;
    // power iteration (models/PointDSC.py:349-356), rows lane and lane+32
    for (int r = lane; r < HY_MAXK; r += 32) V[r] = 1.0f;
    __syncwarp();
    unsigned int notconv = 0u;
    for (int it = 0; it < T; ++it) {
        const float* vin = V + (size_t)it * HY_MAXK;
        float* vout = V + (size_t)(it + 1) * HY_MAXK;
        float y[2] = {0.f, 0.f};
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int r = lane + 32 * h;
            if (r < k) {
                float s = 0.f;
                for (int j = 0; j < k; ++j) s = fmaf(M[r * ldm + j], vin[j], s);
                y[h] = s;
            }
        }
        const float nrm = sqrtf(warp_sum(__fmul_rn(y[0], y[0]) + __fmul_rn(y[1], y[1])));
        bool close = true;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int r = lane + 32 * h;
            if (r < k) {
                const float vn = __fdiv_rn(y[h], nrm + 1e-6f);
                const float vl = vin[r];
                // torch.allclose: |a-b| <= atol + rtol*|b|, rtol=1e-5, atol=1e-8
                close = close && (fabsf(vn - vl) <= 1e-8f + 1e-5f * fabsf(vl));
                vout[r] = vn;
            }
        }
        if (!__all_sync(0xffffffffu, close)) notconv |= 1u << it;
        __syncwarp();
    }
    if (lane == 0 && notconv) atomicOr(notconv_mask + b, notconv);
    // every iterate goes out: which one the reference ends on is only known once all seeds of the pair have voted
    // (seed_kabsch_kernel).  The Kabsch step itself (a serial double-precision 3x3 Jacobi, ~40k clk on one lane) is
    // NOT done here, where it would pin this CTA's 30 KB of shared memory to a single busy lane.
    float* itp = iterates + (size_t)row_id * T * k;
    for (int i = lane; i < T * k; i += 32) itp[i] = V[(size_t)(i / k + 1) * HY_MAXK + (i % k)];
}

// Weighted Kabsch of every seed from the iterate the reference stops on: stop = first iteration at which every seed of
// the pair had converged (models/PointDSC.py:349-356 tests allclose over the whole [S,k] tensor).  One warp per seed,
// HY_KABSCH_WARPS seeds per CTA, so that ~64 of the serial Jacobi solves are in flight per SM.
constexpr int HY_KABSCH_WARPS = 4;
__global__ void __launch_bounds__(32 * HY_KABSCH_WARPS) seed_kabsch_kernel(const float* __restrict__ src,
                                                                           const float* __restrict__ tgt,
                                                                           const int64_t* __restrict__ knn_idx, int N, int S, int k,
                                                                           int T, int n_seeds, const float* __restrict__ iterates,
                                                                           const unsigned int* __restrict__ notconv_mask,
                                                                           float* __restrict__ seed_trans,
                                                                           float* __restrict__ weights,
                                                                           int32_t* __restrict__ iters_out) {
    __shared__ float ps_all[HY_KABSCH_WARPS][HY_MAXK][3], pt_all[HY_KABSCH_WARPS][HY_MAXK][3];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row_id = blockIdx.x * HY_KABSCH_WARPS + warp;
    if (row_id >= n_seeds) return;
    float (*ps)[3] = ps_all[warp];
    float (*pt)[3] = pt_all[warp];
    const int b = row_id / S;
    const unsigned int conv = ~notconv_mask[b] & ((T >= 32) ? 0xffffffffu : ((1u << T) - 1u));
    const int used = conv ? (__ffs(conv)) : T;   // iterations actually run by the reference
    if (iters_out && row_id % S == 0 && lane == 0) iters_out[b] = used;
    for (int r = lane; r < k; r += 32) {
        const int j = (int)knn_idx[(size_t)row_id * k + r];
        for (int d = 0; d < 3; ++d) {
            ps[r][d] = src[((size_t)b * N + j) * 3 + d];
            pt[r][d] = tgt[((size_t)b * N + j) * 3 + d];
        }
    }
    __syncwarp();
    seed_kabsch_warp(iterates + ((size_t)row_id * T + (used - 1)) * k, ps, pt, k, seed_trans + (size_t)row_id * 16,
                     weights ? weights + (size_t)row_id * k : nullptr);
}

size_t seed_hypotheses_workspace(int B, int S, int k, int num_iterations) {
    return align_up(sizeof(unsigned int) * (size_t)B, 256) + sizeof(float) * (size_t)B * S * num_iterations * k;
}

int launch_seed_hypotheses(const float* normed, const float* src, const float* tgt, const int64_t* knn_idx, int B,
                           int N, int C, int S, int k, float sigma, float sigma_spat, int num_iterations,
                           float* seed_trans, float* weights, int32_t* iters, void* ws, size_t ws_bytes,
                           cudaStream_t st) {
    if (k < 1 || k > HY_MAXK) { set_error("seed_hypotheses: k=%d unsupported (1..%d)", k, HY_MAXK); return PDSC_ERR_ARG; }
    if (C % 4 != 0) { set_error("seed_hypotheses: C=%d must be a multiple of 4", C); return PDSC_ERR_ARG; }
    if (num_iterations < 1 || num_iterations > 31) { set_error("seed_hypotheses: num_iterations=%d unsupported (1..31)", num_iterations); return PDSC_ERR_ARG; }
    if (ws_bytes < seed_hypotheses_workspace(B, S, k, num_iterations)) {
        set_error("seed_hypotheses: workspace too small");
        return PDSC_ERR_WORKSPACE;
    }
    unsigned int* mask = reinterpret_cast<unsigned int*>(ws);
    float* iterates = reinterpret_cast<float*>(reinterpret_cast<char*>(ws) + align_up(sizeof(unsigned int) * (size_t)B, 256));
    cudaMemsetAsync(mask, 0, sizeof(unsigned int) * (size_t)B, st);
    const int kp = (k + 3) & ~3;
    const size_t smem = sizeof(float) * ((size_t)kp * (C + 4) + (size_t)kp * (kp + 1) +
                                         (size_t)(num_iterations + 1) * HY_MAXK);
    if (smem > 200 * 1024) { set_error("seed_hypotheses: C=%d too large for shared memory", C); return PDSC_ERR_ARG; }
    cudaFuncSetAttribute(seed_hypo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    // sigma**2 as torch computes it on an fp32 parameter (models/PointDSC.py:259,270)
    const float sig2_feat = sigma * sigma, sig2_spat = sigma_spat * sigma_spat;
    seed_hypo_kernel<<<B * S, HY_THREADS, smem, st>>>(normed, src, tgt, knn_idx, N, C, S, k, sig2_feat, sig2_spat,
                                                      num_iterations, seed_trans, weights, iterates, mask);
    seed_kabsch_kernel<<<(B * S + HY_KABSCH_WARPS - 1) / HY_KABSCH_WARPS, 32 * HY_KABSCH_WARPS, 0, st>>>(
        src, tgt, knn_idx, N, S, k, num_iterations, B * S, iterates, mask, seed_trans, weights, iters);
    count_launches(2);
    return check_launch("seed_hypotheses");
}

}  // namespace pdsc
