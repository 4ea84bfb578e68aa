// Feature head: what PointDSC.forward does with corr_features right after the encoder --
//   normed = F.normalize(corr_features, p=2, dim=1)                   (models/PointDSC.py:156)
//   confidence = classification(corr_features)                        (:107-113, :171; Conv1d 128->32, ReLU, 32->32, ReLU, 32->1)
// -- in ONE pass over the features.  As torch ops this was a norm reduction, a division, a clamp and three small library
// GEMMs with their epilogues (~170 us for 16 x 5000 rows, each re-reading the 41 MB feature matrix); here a CTA stages a
// 128-row tile in shared memory with coalesced loads, threads t and t + 128 walk one half of row t each (sum of squares +
// the 128 -> 32 layer, weights broadcast from shared memory; the halves meet in shared memory), thread t finishes the two
// small layers in registers, and the tile leaves normalised with coalesced stores.  fp32 FMA throughout; the summation
// order differs from cuBLAS, i.e. results agree with the torch path to fp32 rounding (tests: 1e-6 relative).
#include "common.cuh"

#include <math.h>

namespace pdsc {

constexpr int HD_THREADS = 256;          // two threads per row, 64 channels each
constexpr int HD_C = 128, HD_H = 32;
constexpr int HD_PITCH = HD_C + 4;          // floats: row t starts at bank 4t mod 32, so row-per-thread LDS.128 are conflict free

struct HeadParams {
    const float* feat; const float* w1; const float* b1; const float* w2; const float* b2; const float* w3; const float* b3;
    float* normed; float* conf;
    long long rows;
    float eps;
    int32_t* overflow;      // caller-owned flag (pdsc_set_overflow_flag) or null: raised when a confidence is not finite
};

__global__ void __launch_bounds__(HD_THREADS) feature_head_kernel(const HeadParams p) {
    extern __shared__ __align__(16) float hsm[];
    float* xs = hsm;                                   // [128][HD_PITCH]
    float* w1 = xs + 128 * HD_PITCH;                   // [128][32]  (y = x @ w1)
    float* w2 = w1 + HD_C * HD_H;                      // [32][32]
    float* b1 = w2 + HD_H * HD_H;                      // [32]
    float* b2 = b1 + HD_H;                             // [32]
    float* w3 = b2 + HD_H;                             // [32]
    float* den = w3 + HD_H;                            // [128] max(norm, eps) per row
    float* part = den + 128;                           // [128][HD_H + 1] partial sums of the upper channel half
    const int tid = threadIdx.x;
    const long long row0 = (long long)blockIdx.x * 128;
    const int nrows = (int)min(128LL, p.rows - row0);

    {
        const float4* src = reinterpret_cast<const float4*>(p.feat + row0 * HD_C);
#pragma unroll 8
        for (int idx = tid; idx < nrows * 32; idx += HD_THREADS) {
            const int r = idx >> 5, c4 = idx & 31;
            *reinterpret_cast<float4*>(xs + r * HD_PITCH + c4 * 4) = __ldg(src + idx);
        }
        for (int i = tid; i < HD_C * HD_H / 4; i += HD_THREADS)
            reinterpret_cast<float4*>(w1)[i] = __ldg(reinterpret_cast<const float4*>(p.w1) + i);
        for (int i = tid; i < HD_H * HD_H / 4; i += HD_THREADS)
            reinterpret_cast<float4*>(w2)[i] = __ldg(reinterpret_cast<const float4*>(p.w2) + i);
        if (tid < HD_H) { b1[tid] = p.b1[tid]; b2[tid] = p.b2[tid]; w3[tid] = p.w3[tid]; }
    }
    __syncthreads();
    const int rt = tid & 127, half = tid >> 7;
    float h1[HD_H];
    float ss = 0.f;
    if (rt < nrows) {
        const float* xr = xs + rt * HD_PITCH;
#pragma unroll
        for (int o = 0; o < HD_H; ++o) h1[o] = 0.f;
#pragma unroll 2
        for (int c4 = half * (HD_C / 8); c4 < (half + 1) * (HD_C / 8); ++c4) {
            const float4 x4 = *reinterpret_cast<const float4*>(xr + c4 * 4);
            const float xv[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                ss = fmaf(xv[e], xv[e], ss);
                const float4* wrow = reinterpret_cast<const float4*>(w1 + (c4 * 4 + e) * HD_H);
#pragma unroll
                for (int o4 = 0; o4 < HD_H / 4; ++o4) {
                    const float4 w = wrow[o4];
                    h1[o4 * 4 + 0] = fmaf(xv[e], w.x, h1[o4 * 4 + 0]);
                    h1[o4 * 4 + 1] = fmaf(xv[e], w.y, h1[o4 * 4 + 1]);
                    h1[o4 * 4 + 2] = fmaf(xv[e], w.z, h1[o4 * 4 + 2]);
                    h1[o4 * 4 + 3] = fmaf(xv[e], w.w, h1[o4 * 4 + 3]);
                }
            }
        }
        if (half == 1) {
#pragma unroll
            for (int o = 0; o < HD_H; ++o) part[rt * (HD_H + 1) + o] = h1[o];
            part[rt * (HD_H + 1) + HD_H] = ss;
        }
    }
    __syncthreads();
    if (half == 0 && rt < nrows) {
#pragma unroll
        for (int o = 0; o < HD_H; ++o) h1[o] = fmaxf((h1[o] + part[rt * (HD_H + 1) + o]) + b1[o], 0.f);
        ss += part[rt * (HD_H + 1) + HD_H];
        float h2[HD_H];
#pragma unroll
        for (int o = 0; o < HD_H; ++o) h2[o] = 0.f;
#pragma unroll
        for (int j = 0; j < HD_H; ++j) {
            const float4* wrow = reinterpret_cast<const float4*>(w2 + j * HD_H);
#pragma unroll
            for (int o4 = 0; o4 < HD_H / 4; ++o4) {
                const float4 w = wrow[o4];
                h2[o4 * 4 + 0] = fmaf(h1[j], w.x, h2[o4 * 4 + 0]);
                h2[o4 * 4 + 1] = fmaf(h1[j], w.y, h2[o4 * 4 + 1]);
                h2[o4 * 4 + 2] = fmaf(h1[j], w.z, h2[o4 * 4 + 2]);
                h2[o4 * 4 + 3] = fmaf(h1[j], w.w, h2[o4 * 4 + 3]);
            }
        }
        float c = 0.f;
#pragma unroll
        for (int o = 0; o < HD_H; ++o) c = fmaf(fmaxf(h2[o] + b2[o], 0.f), w3[o], c);
        const float cv = c + p.b3[0];
        p.conf[row0 + rt] = cv;
        if (p.overflow && !(fabsf(cv) <= 3.4e38f)) atomicOr(p.overflow, 1);     // inf or NaN (an upstream overflow that survived)
        den[rt] = fmaxf(sqrtf(ss), p.eps);            // F.normalize: x / max(||x||_2, eps)
    }
    __syncthreads();
    {
        float4* dst = reinterpret_cast<float4*>(p.normed + row0 * HD_C);
#pragma unroll 8
        for (int idx = tid; idx < nrows * 32; idx += HD_THREADS) {
            const int r = idx >> 5, c4 = idx & 31;
            const float4 v = *reinterpret_cast<const float4*>(xs + r * HD_PITCH + c4 * 4);
            const float d = den[r];
            dst[idx] = make_float4(__fdiv_rn(v.x, d), __fdiv_rn(v.y, d), __fdiv_rn(v.z, d), __fdiv_rn(v.w, d));
        }
    }
}

int launch_feature_head(const float* feat, long long rows, int C, int H1, int H2, const float* w1t, const float* b1, const float* w2t,
                        const float* b2, const float* w3, const float* b3, float eps, float* normed, float* conf, cudaStream_t st) {
    if (C != HD_C || H1 != HD_H || H2 != HD_H) {
        set_error("feature_head: built for C=%d and a %d-%d-1 classifier (got C=%d, %d-%d-1)", HD_C, HD_H, HD_H, C, H1, H2);
        return PDSC_ERR_ARG;
    }
    HeadParams p;
    p.feat = feat; p.w1 = w1t; p.b1 = b1; p.w2 = w2t; p.b2 = b2; p.w3 = w3; p.b3 = b3;
    p.normed = normed; p.conf = conf; p.rows = rows; p.eps = eps;
    p.overflow = overflow_flag();
    const int smem = (int)sizeof(float) * (128 * HD_PITCH + HD_C * HD_H + HD_H * HD_H + 3 * HD_H + 128 + 128 * (HD_H + 1));
    opt_in_smem<feature_head_kernel>(smem);
    feature_head_kernel<<<(unsigned)((rows + 127) / 128), HD_THREADS, smem, st>>>(p);
    count_launches(1);
    return check_launch("feature_head");
}

}  // namespace pdsc
