// (a) Compatibility-weighted non-local attention, fp32 CUDA-core flavour (PDSC_ATTN_FP32).
// Replaces models/PointDSC.py:150-153 + NonLocalBlock.forward :39-42.  Flash-style: the N x N
// compatibility matrix, the logits and the softmax weights live only in registers / shared
// memory; HBM sees Q,K,V [N,C] in and the message [N,C] out.
// This flavour does the two contractions with fp32 FMAs in the reference's own arithmetic
// (IEEE sqrt and divide in the compatibility term, expf); it is the exact-parity mode and the
// yardstick for the tcgen05 kernel in attn_tc.cu.
#include "common.cuh"

#include <math.h>

namespace pdsc {

constexpr int AF_C = 128;
constexpr int AF_BM = 64, AF_BN = 64, AF_THREADS = 256;
constexpr int AF_LD = AF_C + 4;     // feature row stride in smem (floats): odd multiple of 4
constexpr int AF_LDP = AF_BN + 4;

struct AttnFp32Smem {
    float Q[AF_BM][AF_LD];
    float K[AF_BN][AF_LD];
    float V[AF_BN][AF_LD];
    float P[AF_BM][AF_LDP];
    float kpt[AF_BN][8];   // src xyz, tgt xyz of the key tile
};

__global__ void __launch_bounds__(AF_THREADS, 1) attn_fp32_kernel(
    const float* __restrict__ q, const float* __restrict__ k, const float* __restrict__ v, int64_t ld,
    const float* __restrict__ src, const float* __restrict__ tgt, float* __restrict__ out, int64_t ld_out, int N,
    float sig2, float scale) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    AttnFp32Smem& sm = *reinterpret_cast<AttnFp32Smem*>(smem_raw);
    const int b = blockIdx.y;
    const int m0 = blockIdx.x * AF_BM;
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const float* qb = q + (size_t)b * N * ld;
    const float* kb = k + (size_t)b * N * ld;
    const float* vb = v + (size_t)b * N * ld;
    const float* sb = src + (size_t)b * N * 3;
    const float* tb = tgt + (size_t)b * N * 3;

    // Q tile (rows past N are zero)
    for (int f = tid; f < AF_BM * (AF_C / 4); f += AF_THREADS) {
        const int r = f / (AF_C / 4), c4 = f % (AF_C / 4);
        float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
        if (m0 + r < N) val = *reinterpret_cast<const float4*>(qb + (size_t)(m0 + r) * ld + c4 * 4);
        *reinterpret_cast<float4*>(&sm.Q[r][c4 * 4]) = val;
    }
    // my 4 query rows: ty + 16 i
    float qs[4][3], qt[4][3];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = min(m0 + ty + 16 * i, N - 1);
#pragma unroll
        for (int d = 0; d < 3; ++d) { qs[i][d] = sb[m * 3 + d]; qt[i][d] = tb[m * 3 + d]; }
    }
    float row_max[4], row_sum[4], o[4][8];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        row_max[i] = -INFINITY;
        row_sum[i] = 0.f;
#pragma unroll
        for (int c = 0; c < 8; ++c) o[i][c] = 0.f;
    }

    for (int n0 = 0; n0 < N; n0 += AF_BN) {
        __syncthreads();   // previous tile fully consumed (also covers the Q fill)
        for (int f = tid; f < AF_BN * (AF_C / 4); f += AF_THREADS) {
            const int r = f / (AF_C / 4), c4 = f % (AF_C / 4);
            float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
            if (n0 + r < N) {
                kv = *reinterpret_cast<const float4*>(kb + (size_t)(n0 + r) * ld + c4 * 4);
                vv = *reinterpret_cast<const float4*>(vb + (size_t)(n0 + r) * ld + c4 * 4);
            }
            *reinterpret_cast<float4*>(&sm.K[r][c4 * 4]) = kv;
            *reinterpret_cast<float4*>(&sm.V[r][c4 * 4]) = vv;
        }
        if (tid < AF_BN) {
            const int n = min(n0 + tid, N - 1);
#pragma unroll
            for (int d = 0; d < 3; ++d) { sm.kpt[tid][d] = sb[n * 3 + d]; sm.kpt[tid][4 + d] = tb[n * 3 + d]; }
        }
        __syncthreads();

        // S = Q K^T for rows ty+16i, keys tx+16j
        float s[4][4] = {};
#pragma unroll 4
        for (int c = 0; c < AF_C; c += 4) {
            float4 qa[4], ka[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                qa[i] = *reinterpret_cast<const float4*>(&sm.Q[ty + 16 * i][c]);
                ka[i] = *reinterpret_cast<const float4*>(&sm.K[tx + 16 * i][c]);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    s[i][j] = fmaf(qa[i].x, ka[j].x, s[i][j]);
                    s[i][j] = fmaf(qa[i].y, ka[j].y, s[i][j]);
                    s[i][j] = fmaf(qa[i].z, ka[j].z, s[i][j]);
                    s[i][j] = fmaf(qa[i].w, ka[j].w, s[i][j]);
                }
        }
        // x = compat * logits, online softmax
        float tile_max[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) tile_max[i] = -INFINITY;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int nl = tx + 16 * j;
            const bool valid = n0 + nl < N;
            const float kx = sm.kpt[nl][0], ky = sm.kpt[nl][1], kz = sm.kpt[nl][2];
            const float ux = sm.kpt[nl][4], uy = sm.kpt[nl][5], uz = sm.kpt[nl][6];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float ds = dist3(qs[i][0], qs[i][1], qs[i][2], kx, ky, kz);
                const float dt = dist3(qt[i][0], qt[i][1], qt[i][2], ux, uy, uz);
                const float x = valid ? __fmul_rn(compat_exact(ds, dt, sig2), __fmul_rn(s[i][j], scale)) : -INFINITY;
                s[i][j] = x;
                tile_max[i] = fmaxf(tile_max[i], x);
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float tm = tile_max[i];
#pragma unroll
            for (int off = 8; off > 0; off >>= 1) tm = fmaxf(tm, __shfl_xor_sync(0xffffffffu, tm, off));
            const float new_max = fmaxf(row_max[i], tm);
            const float corr = expf(row_max[i] - new_max);   // exp(-inf) = 0 on the first tile
            float psum = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float p = expf(s[i][j] - new_max);
                psum += p;
                sm.P[ty + 16 * i][tx + 16 * j] = p;
            }
#pragma unroll
            for (int off = 8; off > 0; off >>= 1) psum += __shfl_xor_sync(0xffffffffu, psum, off);
            row_sum[i] = row_sum[i] * corr + psum;
            row_max[i] = new_max;
#pragma unroll
            for (int c = 0; c < 8; ++c) o[i][c] *= corr;
        }
        __syncwarp();   // P rows ty+16i are produced and consumed by the same half warp
        // O += P V, channels tx*4..+3 and 64+tx*4..+3
#pragma unroll 2
        for (int n = 0; n < AF_BN; n += 4) {
            float4 pa[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) pa[i] = *reinterpret_cast<const float4*>(&sm.P[ty + 16 * i][n]);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float4 v0 = *reinterpret_cast<const float4*>(&sm.V[n + e][tx * 4]);
                const float4 v1 = *reinterpret_cast<const float4*>(&sm.V[n + e][64 + tx * 4]);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float p = e == 0 ? pa[i].x : (e == 1 ? pa[i].y : (e == 2 ? pa[i].z : pa[i].w));
                    o[i][0] = fmaf(p, v0.x, o[i][0]); o[i][1] = fmaf(p, v0.y, o[i][1]);
                    o[i][2] = fmaf(p, v0.z, o[i][2]); o[i][3] = fmaf(p, v0.w, o[i][3]);
                    o[i][4] = fmaf(p, v1.x, o[i][4]); o[i][5] = fmaf(p, v1.y, o[i][5]);
                    o[i][6] = fmaf(p, v1.z, o[i][6]); o[i][7] = fmaf(p, v1.w, o[i][7]);
                }
            }
        }
    }
    float* ob = out + (size_t)b * N * ld_out;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = m0 + ty + 16 * i;
        if (m >= N) continue;
        const float inv = 1.0f / row_sum[i];
        *reinterpret_cast<float4*>(ob + (size_t)m * ld_out + tx * 4) =
            make_float4(o[i][0] * inv, o[i][1] * inv, o[i][2] * inv, o[i][3] * inv);
        *reinterpret_cast<float4*>(ob + (size_t)m * ld_out + 64 + tx * 4) =
            make_float4(o[i][4] * inv, o[i][5] * inv, o[i][6] * inv, o[i][7] * inv);
    }
}

int launch_attn_fp32(const float* q, const float* k, const float* v, int64_t ld, const float* src, const float* tgt,
                     float* out, int64_t ld_out, int B, int N, int C, float sigma_spat, cudaStream_t st) {
    if (C != AF_C) { set_error("nonlocal_attn: C=%d unsupported (kernels are built for C=%d)", C, AF_C); return PDSC_ERR_ARG; }
    opt_in_smem<attn_fp32_kernel>((int)sizeof(AttnFp32Smem));
    const float scale = 1.0f / sqrtf((float)C);
    dim3 grid((N + AF_BM - 1) / AF_BM, B);
    attn_fp32_kernel<<<grid, AF_THREADS, sizeof(AttnFp32Smem), st>>>(q, k, v, ld, src, tgt, out, ld_out, N,
                                                                     sigma_spat * sigma_spat, scale);
    count_launches(1);
    return check_launch("nonlocal_attn(fp32)");
}

}  // namespace pdsc
