// (e) Hypothesis scoring: every seed transform against all N correspondences.
// Replaces models/PointDSC.py:325-335 (einsum -> [1,S,N,3] temp, norm, compare, mean, argmax,
// two gathers).  Integer inlier counts are produced directly; nothing of size S*N touches HBM.
#include "common.cuh"

namespace pdsc {

constexpr int SC_THREADS = 128;   // seeds per block (one seed per thread)
constexpr int SC_POINTS = 512;    // correspondences staged in shared memory per block

__device__ __forceinline__ bool is_inlier(const float* T, float sx, float sy, float sz, float tx, float ty, float tz,
                                          float thr) {
    float px = fmaf(T[2], sz, fmaf(T[1], sy, T[0] * sx)) + T[3];
    float py = fmaf(T[6], sz, fmaf(T[5], sy, T[4] * sx)) + T[7];
    float pz = fmaf(T[10], sz, fmaf(T[9], sy, T[8] * sx)) + T[11];
    return dist3(px, py, pz, tx, ty, tz) < thr;
}
// Same decision without the square root: thr2 = squared_threshold(thr) (common.cuh), d2 rounded like dist3's argument.
__device__ __forceinline__ bool is_inlier_sq(const float* T, float sx, float sy, float sz, float tx, float ty, float tz,
                                             float thr2) {
    float px = fmaf(T[2], sz, fmaf(T[1], sy, T[0] * sx)) + T[3];
    float py = fmaf(T[6], sz, fmaf(T[5], sy, T[4] * sx)) + T[7];
    float pz = fmaf(T[10], sz, fmaf(T[9], sy, T[8] * sx)) + T[11];
    const float dx = px - tx, dy = py - ty, dz = pz - tz;
    const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
    return d2 < thr2;
}

// Two correspondences per packed f32x2 instruction (each lane rounded like the scalar operation of is_inlier_sq: the
// counts are identical).  The two sums of the squared distance stay scalar: ptxas contracts mul.rn.f32x2 + add.rn.f32x2.
__global__ void __launch_bounds__(SC_THREADS) score_count_kernel(const float* __restrict__ seed_trans,
                                                                 const float* __restrict__ src,
                                                                 const float* __restrict__ tgt, int N, int S,
                                                                 float thr2, int32_t* __restrict__ inlier_count) {
    // per pair of points: (sx0, sx1, sy0, sy1), (sz0, sz1, -tx0, -tx1), (-ty0, -ty1, -tz0, -tz1)
    __shared__ float4 sp[SC_POINTS / 2 * 3];
    const int b = blockIdx.z;
    const int n0 = blockIdx.y * SC_POINTS;
    const int np = min(SC_POINTS, N - n0);
    const int np2 = (np + 1) >> 1;
    const float* sb = src + ((size_t)b * N + n0) * 3;
    const float* tb = tgt + ((size_t)b * N + n0) * 3;
    for (int i = threadIdx.x; i < np2; i += SC_THREADS) {
        const int n = 2 * i;
        const bool two = n + 1 < np;
        // an odd tail is padded with a target infinitely far away: never an inlier
        const float sx1 = two ? sb[n * 3 + 3] : 0.f, sy1 = two ? sb[n * 3 + 4] : 0.f, sz1 = two ? sb[n * 3 + 5] : 0.f;
        const float tx1 = two ? tb[n * 3 + 3] : INFINITY, ty1 = two ? tb[n * 3 + 4] : INFINITY, tz1 = two ? tb[n * 3 + 5] : INFINITY;
        sp[i * 3 + 0] = make_float4(sb[n * 3], sx1, sb[n * 3 + 1], sy1);
        sp[i * 3 + 1] = make_float4(sb[n * 3 + 2], sz1, -tb[n * 3], -tx1);
        sp[i * 3 + 2] = make_float4(-tb[n * 3 + 1], -ty1, -tb[n * 3 + 2], -tz1);
    }
    __syncthreads();
    const int s = blockIdx.x * SC_THREADS + threadIdx.x;
    if (s >= S) return;
    float2 T[12];
    const float4* tp = reinterpret_cast<const float4*>(seed_trans + ((size_t)b * S + s) * 16);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const float4 row = tp[r];
        T[r * 4 + 0] = make_float2(row.x, row.x); T[r * 4 + 1] = make_float2(row.y, row.y);
        T[r * 4 + 2] = make_float2(row.z, row.z); T[r * 4 + 3] = make_float2(row.w, row.w);
    }
    int cnt = 0;
#pragma unroll 4
    for (int i = 0; i < np2; ++i) {
        const float4 a = sp[i * 3], c = sp[i * 3 + 1], e = sp[i * 3 + 2];
        const float2 sx = make_float2(a.x, a.y), sy = make_float2(a.z, a.w), sz = make_float2(c.x, c.y);
        // p = fma(T2, sz, fma(T1, sy, T0 * sx)) + T3, exactly as is_inlier_sq
        const float2 px = __fadd2_rn(__ffma2_rn(T[2], sz, __ffma2_rn(T[1], sy, __fmul2_rn(T[0], sx))), T[3]);
        const float2 py = __fadd2_rn(__ffma2_rn(T[6], sz, __ffma2_rn(T[5], sy, __fmul2_rn(T[4], sx))), T[7]);
        const float2 pz = __fadd2_rn(__ffma2_rn(T[10], sz, __ffma2_rn(T[9], sy, __fmul2_rn(T[8], sx))), T[11]);
        const float2 dx = __fadd2_rn(px, make_float2(c.z, c.w));
        const float2 dy = __fadd2_rn(py, make_float2(e.x, e.y));
        const float2 dz = __fadd2_rn(pz, make_float2(e.z, e.w));
        const float2 qx = __fmul2_rn(dx, dx), qy = __fmul2_rn(dy, dy), qz = __fmul2_rn(dz, dz);
        const float d0 = __fadd_rn(__fadd_rn(qx.x, qy.x), qz.x), d1 = __fadd_rn(__fadd_rn(qx.y, qy.y), qz.y);
        cnt += (d0 < thr2 ? 1 : 0) + (d1 < thr2 ? 1 : 0);
    }
    atomicAdd(inlier_count + (size_t)b * S + s, cnt);
}

// One block per pair: first-index argmax of the counts (torch.argmax on CPU returns the first
// maximal index; fitness = count/N is strictly monotone in count for N < 2^23), then the best
// transform and the 0/1 labels under it (models/PointDSC.py:330-335).
__global__ void __launch_bounds__(256) score_select_kernel(const float* __restrict__ seed_trans,
                                                           const float* __restrict__ src,
                                                           const float* __restrict__ tgt, int N, int S, float thr,
                                                           const int32_t* __restrict__ inlier_count,
                                                           int64_t* __restrict__ best_idx,
                                                           float* __restrict__ final_trans,
                                                           float* __restrict__ final_labels) {
    __shared__ unsigned long long red[8];
    __shared__ float Tb[16];
    const int b = blockIdx.x;
    // key = count << 32 | (0xffffffff - idx): max key == max count, then min index
    unsigned long long best = 0ull;
    for (int s = threadIdx.x; s < S; s += blockDim.x) {
        unsigned long long key = ((unsigned long long)(uint32_t)inlier_count[(size_t)b * S + s] << 32) |
                                 (unsigned long long)(0xffffffffu - (uint32_t)s);
        best = key > best ? key : best;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
        best = other > best ? other : best;
    }
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) best = red[w] > best ? red[w] : best;
        red[0] = best;
    }
    __syncthreads();
    const int sbest = (int)(0xffffffffu - (uint32_t)(red[0] & 0xffffffffull));
    if (threadIdx.x < 16) {
        float v = seed_trans[((size_t)b * S + sbest) * 16 + threadIdx.x];
        Tb[threadIdx.x] = v;
        final_trans[(size_t)b * 16 + threadIdx.x] = v;
    }
    if (threadIdx.x == 0) best_idx[b] = sbest;
    __syncthreads();
    const float* sb = src + (size_t)b * N * 3;
    const float* tb = tgt + (size_t)b * N * 3;
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
        bool in = is_inlier(Tb, sb[n * 3], sb[n * 3 + 1], sb[n * 3 + 2], tb[n * 3], tb[n * 3 + 1], tb[n * 3 + 2], thr);
        final_labels[(size_t)b * N + n] = in ? 1.0f : 0.0f;
    }
}

int launch_score_hypotheses(const float* seed_trans, const float* src, const float* tgt, int B, int N, int S,
                            float thr, int32_t* inlier_count, int64_t* best_idx, float* final_trans,
                            float* final_labels, cudaStream_t st) {
    cudaMemsetAsync(inlier_count, 0, sizeof(int32_t) * (size_t)B * S, st);
    dim3 grid((S + SC_THREADS - 1) / SC_THREADS, (N + SC_POINTS - 1) / SC_POINTS, B);
    score_count_kernel<<<grid, SC_THREADS, 0, st>>>(seed_trans, src, tgt, N, S, squared_threshold(thr), inlier_count);
    score_select_kernel<<<B, 256, 0, st>>>(seed_trans, src, tgt, N, S, thr, inlier_count, best_idx, final_trans,
                                           final_labels);
    count_launches(2);
    return check_launch("score_hypotheses");
}

}  // namespace pdsc
