// (e) Post refinement: PointDSC.post_refinement (models/PointDSC.py:403-438) as ONE kernel.
// One CTA per pair runs the whole <=20 round loop on the device: transform, distance, mask,
// inlier count, the "count repeated -> stop" test, weighted Kabsch (two-pass centroid form of
// models/common.py:24-33) and the 3x3 SVD.  The reference does ~15 launches, a diag_embed of
// [1,n,n] and two host syncs (int(), svd on CPU) per round.
#include "common.cuh"

namespace pdsc {

constexpr int RF_THREADS = 1024;

template <int NV>
__device__ __forceinline__ void block_reduce(double (&v)[NV], double* smem /* [32*NV] */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < NV; ++i) v[i] = warp_sum(v[i]);
    __syncthreads();  // smem reuse across calls
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < NV; ++i) smem[warp * NV + i] = v[i];
    }
    __syncthreads();
    const int nw = blockDim.x >> 5;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        double x = (lane < nw) ? smem[lane * NV + i] : 0.0;
        v[i] = warp_sum(x);
    }
}

__global__ void __launch_bounds__(RF_THREADS) post_refine_kernel(float* __restrict__ trans,
                                                                 const float* __restrict__ src,
                                                                 const float* __restrict__ tgt, int N, float tau,
                                                                 int max_iter, int32_t* __restrict__ iters_out,
                                                                 int32_t* __restrict__ counts_out) {
    __shared__ double red[32 * 9];
    __shared__ float T[16];
    const int b = blockIdx.x;
    const float* sb = src + (size_t)b * N * 3;
    const float* tb = tgt + (size_t)b * N * 3;
    if (threadIdx.x < 16) T[threadIdx.x] = trans[(size_t)b * 16 + threadIdx.x];
    if (counts_out)
        for (int i = threadIdx.x; i < max_iter; i += blockDim.x) counts_out[(size_t)b * max_iter + i] = -1;
    __syncthreads();

    long long prev = 0;
    int refits = 0;
    for (int it = 0; it < max_iter; ++it) {
        // pass 1: mask, weights, weighted sums
        double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int n = threadIdx.x; n < N; n += blockDim.x) {
            float ax = sb[n * 3], ay = sb[n * 3 + 1], az = sb[n * 3 + 2];
            float bx = tb[n * 3], by = tb[n * 3 + 1], bz = tb[n * 3 + 2];
            float px = fmaf(T[2], az, fmaf(T[1], ay, T[0] * ax)) + T[3];
            float py = fmaf(T[6], az, fmaf(T[5], ay, T[4] * ax)) + T[7];
            float pz = fmaf(T[10], az, fmaf(T[9], ay, T[8] * ax)) + T[11];
            float d = dist3(px, py, pz, bx, by, bz);
            if (d < tau) {
                float r = __fdiv_rn(d, tau);
                float w = __fdiv_rn(1.0f, __fadd_rn(1.0f, __fmul_rn(r, r)));
                acc[0] += 1.0;
                acc[1] += (double)w;
                acc[2] += (double)__fmul_rn(ax, w); acc[3] += (double)__fmul_rn(ay, w); acc[4] += (double)__fmul_rn(az, w);
                acc[5] += (double)__fmul_rn(bx, w); acc[6] += (double)__fmul_rn(by, w); acc[7] += (double)__fmul_rn(bz, w);
            }
        }
        block_reduce<8>(acc, red);
        const long long cnt = (long long)(acc[0] + 0.5);
        if (threadIdx.x == 0 && counts_out) counts_out[(size_t)b * max_iter + it] = (int32_t)cnt;
        if (cnt == prev) break;   // abs(int(inlier_num - previous_inlier_num)) < 1
        prev = cnt;
        const float wsum = (float)acc[1] + 1e-6f;
        const float cA[3] = {__fdiv_rn((float)acc[2], wsum), __fdiv_rn((float)acc[3], wsum), __fdiv_rn((float)acc[4], wsum)};
        const float cB[3] = {__fdiv_rn((float)acc[5], wsum), __fdiv_rn((float)acc[6], wsum), __fdiv_rn((float)acc[7], wsum)};
        // pass 2: H = (A-cA)^T diag(w) (B-cB)
        double h[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        for (int n = threadIdx.x; n < N; n += blockDim.x) {
            float ax = sb[n * 3], ay = sb[n * 3 + 1], az = sb[n * 3 + 2];
            float bx = tb[n * 3], by = tb[n * 3 + 1], bz = tb[n * 3 + 2];
            float px = fmaf(T[2], az, fmaf(T[1], ay, T[0] * ax)) + T[3];
            float py = fmaf(T[6], az, fmaf(T[5], ay, T[4] * ax)) + T[7];
            float pz = fmaf(T[10], az, fmaf(T[9], ay, T[8] * ax)) + T[11];
            float d = dist3(px, py, pz, bx, by, bz);
            if (d < tau) {
                float r = __fdiv_rn(d, tau);
                double w = (double)__fdiv_rn(1.0f, __fadd_rn(1.0f, __fmul_rn(r, r)));
                double am[3] = {(double)(ax - cA[0]), (double)(ay - cA[1]), (double)(az - cA[2])};
                double bm[3] = {(double)(bx - cB[0]), (double)(by - cB[1]), (double)(bz - cB[2])};
#pragma unroll
                for (int i = 0; i < 3; ++i)
#pragma unroll
                    for (int j = 0; j < 3; ++j) h[i * 3 + j] += w * am[i] * bm[j];
            }
        }
        block_reduce<9>(h, red);
        __syncthreads();   // everyone is done reading T
        if (threadIdx.x == 0) {
            double R[9];
            kabsch_rotation(h, R);
            write_transform(T, R, cA, cB);
        }
        ++refits;
        __syncthreads();
    }
    if (threadIdx.x < 16) trans[(size_t)b * 16 + threadIdx.x] = T[threadIdx.x];
    if (threadIdx.x == 0 && iters_out) iters_out[b] = refits;
}

int launch_post_refine(float* trans, const float* src, const float* tgt, int B, int N, float tau, int max_iter,
                       int32_t* iters_out, int32_t* counts_out, cudaStream_t st) {
    post_refine_kernel<<<B, RF_THREADS, 0, st>>>(trans, src, tgt, N, tau, max_iter, iters_out, counts_out);
    count_launches(1);
    return check_launch("post_refine");
}

}  // namespace pdsc
