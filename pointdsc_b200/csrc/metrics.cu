// (N3) Evaluation metrics on the device, the step right after PointDSC.forward.
// Replaces the per-pair Python loops (and the sklearn / CPU round trip) of
//   libs/loss.py:34-63   TransformationLoss.forward: RE [deg], TE [cm], RMSE, squared-residual loss, recall flag
//   libs/loss.py:96-112  ClassificationLoss.forward: precision / recall / F1 counts, mean logit of inliers / outliers
// One CTA per registration pair; the reductions over the N correspondences are integer counts (bit-exact) and
// fp32 sums accumulated in double (the reference sums in fp32; the difference is below its own rounding noise).
#include "common.cuh"

namespace pdsc {

constexpr int MT_THREADS = 256;

template <typename T>
__device__ T block_sum(T v, T* scratch) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
    __syncthreads();
    T t = 0;
    for (int w = 0; w < MT_THREADS / 32; ++w) t += scratch[w];
    return t;
}

// out[p] = {re_deg, te_cm, rmse, mean squared residual}; trans / gt_trans row-major 4x4
__global__ void __launch_bounds__(MT_THREADS) registration_metrics_kernel(const float* __restrict__ trans, const float* __restrict__ gt,
                                                                          const float* __restrict__ src, const float* __restrict__ tgt,
                                                                          int N, float* __restrict__ out) {
    __shared__ double scratch[MT_THREADS / 32];
    const int p = blockIdx.x;
    const float* T = trans + (size_t)p * 16;
    const float* G = gt + (size_t)p * 16;
    src += (size_t)p * N * 3;
    tgt += (size_t)p * N * 3;
    float R[9], t[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
        for (int c = 0; c < 3; ++c) R[r * 3 + c] = T[r * 4 + c];
        t[r] = T[r * 4 + 3];
    }
    double s_norm = 0.0, s_sq = 0.0;
    for (int n = threadIdx.x; n < N; n += MT_THREADS) {
        const float x = src[3 * n], y = src[3 * n + 1], z = src[3 * n + 2];
        // transform(): R @ pts^T + t  (utils/SE3.py:43-57)
        const float wx = fmaf(R[2], z, fmaf(R[1], y, R[0] * x)) + t[0];
        const float wy = fmaf(R[5], z, fmaf(R[4], y, R[3] * x)) + t[1];
        const float wz = fmaf(R[8], z, fmaf(R[7], y, R[6] * x)) + t[2];
        const float dx = wx - tgt[3 * n], dy = wy - tgt[3 * n + 1], dz = wz - tgt[3 * n + 2];
        const float sq = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
        s_sq += (double)sq;
        s_norm += (double)sqrtf(sq);
    }
    s_norm = block_sum(s_norm, scratch);
    s_sq = block_sum(s_sq, scratch);
    if (threadIdx.x == 0) {
        // trace(R^T R_gt) = sum_ij R_ij Rgt_ij
        float tr = 0.f;
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c) tr = fmaf(R[r * 3 + c], G[r * 4 + c], tr);
        const float cosv = fminf(fmaxf((tr - 1.0f) * 0.5f, -1.0f), 1.0f);
        const float re = acosf(cosv) * 180.0f / 3.14159265358979323846f;
        const float ex = t[0] - G[3], ey = t[1] - G[7], ez = t[2] - G[11];
        const float te = sqrtf(ex * ex + ey * ey + ez * ez) * 100.0f;
        out[p * 4 + 0] = re;
        out[p * 4 + 1] = te;
        out[p * 4 + 2] = (float)(s_norm / N);
        out[p * 4 + 3] = (float)(s_sq / N);
    }
}

// counts[p] = {tp, fp, fn, tn} with predicted label = pred > thr, ground truth = gt != 0;
// sums[p] = {sum pred*gt, sum pred*(1-gt)}
__global__ void __launch_bounds__(MT_THREADS) classification_counts_kernel(const float* __restrict__ pred, const float* __restrict__ gt, int N,
                                                                           float thr, int32_t* __restrict__ counts, float* __restrict__ sums) {
    __shared__ double dscratch[MT_THREADS / 32];
    __shared__ int iscratch[MT_THREADS / 32];
    const int p = blockIdx.x;
    pred += (size_t)p * N;
    gt += (size_t)p * N;
    int tp = 0, fp = 0, fn = 0, tn = 0;
    double st = 0.0, sf = 0.0;
    for (int n = threadIdx.x; n < N; n += MT_THREADS) {
        const float x = pred[n], g = gt[n];
        const bool pl = x > thr, gl = g != 0.f;
        tp += pl && gl; fp += pl && !gl; fn += !pl && gl; tn += !pl && !gl;
        st += (double)(x * g);
        sf += (double)(x * (1.0f - g));
    }
    tp = block_sum(tp, iscratch); fp = block_sum(fp, iscratch); fn = block_sum(fn, iscratch); tn = block_sum(tn, iscratch);
    st = block_sum(st, dscratch); sf = block_sum(sf, dscratch);
    if (threadIdx.x == 0) {
        counts[p * 4 + 0] = tp; counts[p * 4 + 1] = fp; counts[p * 4 + 2] = fn; counts[p * 4 + 3] = tn;
        sums[p * 2 + 0] = (float)st; sums[p * 2 + 1] = (float)sf;
    }
}

int launch_registration_metrics(const float* trans, const float* gt, const float* src, const float* tgt, int P, int N, float* out,
                                cudaStream_t st) {
    registration_metrics_kernel<<<P, MT_THREADS, 0, st>>>(trans, gt, src, tgt, N, out);
    count_launches(1);
    return check_launch("registration_metrics");
}

int launch_classification_counts(const float* pred, const float* gt, int P, int N, float thr, int32_t* counts, float* sums,
                                 cudaStream_t st) {
    classification_counts_kernel<<<P, MT_THREADS, 0, st>>>(pred, gt, N, thr, counts, sums);
    count_launches(1);
    return check_launch("classification_counts");
}

}  // namespace pdsc
