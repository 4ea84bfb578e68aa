// Shared device/host helpers for the pointdsc_b200 kernels (sm_100a).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/pointdsc_b200.h"

namespace pdsc {

// ---- host side error plumbing (thread local, read through pdsc_last_error) -----------------
void set_error(const char* fmt, ...);
int check_launch(const char* what);
void count_launches(int n);   // kernels enqueued, read back through pdsc_launch_count()

#define PDSC_REQUIRE(cond, ...)          \
    do {                                 \
        if (!(cond)) {                   \
            pdsc::set_error(__VA_ARGS__); \
            return PDSC_ERR_ARG;         \
        }                                \
    } while (0)

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
int num_sms();

// Opt-in to more than 48 KB of dynamic shared memory.  The attribute is per device (and per function), so the cache is
// keyed by the current device: a process that runs on cuda:0 and then on cuda:1 opts in on both.
// (`static`: internal linkage.  An inline template's local static is a GNU-unique symbol, ONE object per process even across
// dlopen(RTLD_LOCAL), so the product and the tuning library -- both loaded by the tests -- would share the cache while each
// has its own kernels.)
template <auto Kernel>
static void opt_in_smem(int bytes) {
    static thread_local int set_bytes[64];      // zero-initialised; per kernel instantiation
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && set_bytes[dev] >= bytes) return;
    cudaFuncSetAttribute(Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (dev >= 0 && dev < 64) set_bytes[dev] = bytes;
}

// ---- operand range of the fp16 hi/lo split ------------------------------------------------------
// A value at or beyond the fp16 range (65504) splits into hi = inf, lo = -inf/NaN; the MMA then produces NaN and the next
// ReLU turns that NaN into 0: silently wrong numbers.  Every split therefore ORs an "exponent all ones" test of its hi word
// into a register and, at the end of the kernel, into the caller's flag (pdsc_set_overflow_flag; null = no flag).
int32_t* overflow_flag();             // the flag registered for the current host thread (device memory owned by the caller)
__device__ __forceinline__ uint32_t half2_not_finite(uint32_t h2) {     // non-zero if either half is inf or NaN
    return ((h2 & 0x7c007c00u) + 0x04000400u) & 0x80008000u;
}

// ---- tuning / debugging knobs ----------------------------------------------------------------
// The shipped library (libpointdsc_b200.so) is built WITHOUT PDSC_TUNING: every knob is its compile-time default and the
// environment is never read.  The tests and scripts/ that exercise other launch shapes load libpointdsc_b200_tuning.so,
// the same sources built with -DPDSC_TUNING, where `tuning_int("PDSC_X", d)` reads the environment variable per call.
#ifdef PDSC_TUNING
int tuning_int(const char* name, int dflt);
const char* tuning_str(const char* name);
constexpr bool kTuningBuild = true;
#else
inline int tuning_int(const char*, int dflt) { return dflt; }
inline const char* tuning_str(const char*) { return nullptr; }
constexpr bool kTuningBuild = false;
#endif

// ---- warp helpers ---------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Monotone map float -> uint32 (ascending float order == ascending unsigned order; -0 < +0).
__device__ __forceinline__ uint32_t float_to_ordered(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// The reference compares sqrt(dx^2+dy^2+dz^2) with a threshold (torch.norm then < or >=).
// Products and sums are rounded separately (no FMA contraction) like a scalar fp32 loop.
__device__ __forceinline__ float dist3(float ax, float ay, float az, float bx, float by, float bz) {
    float dx = ax - bx, dy = ay - by, dz = az - bz;
    float s = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
    return sqrtf(s);
}

// Length compatibility of correspondences i,j (models/PointDSC.py:151-153):
// max(0, 1 - (|si-sj| - |ti-tj|)^2 / sigma^2), sig2 = sigma*sigma rounded to fp32 like torch.
__device__ __forceinline__ float compat_exact(float ds, float dt, float sig2) {
    float c = ds - dt;
    return fmaxf(__fsub_rn(1.0f, __fdiv_rn(__fmul_rn(c, c), sig2)), 0.0f);
}

// ---- 3x3 Kabsch rotation ---------------------------------------------------------------------
// R = V diag(1,1,det(V U^T)) U^T for H = U S V^T (models/common.py:35-41).  That product equals
// [v1 v2 v1xv2][u1 u2 u1xu2]^T whatever signs LAPACK picks, so only the two leading singular
// pairs are needed.  Done in double: eigenvectors of H^T H by cyclic Jacobi, u_i = H v_i/|H v_i|.
// flag: 0 ok, 1 rank<=1 (u2 arbitrary), 2 H == 0 (identity returned).
__device__ inline int kabsch_rotation(const double H[9], double R[9]) {
    // A = H^T H (symmetric)
    double A[3][3], V[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            A[i][j] = H[0 * 3 + i] * H[0 * 3 + j] + H[1 * 3 + i] * H[1 * 3 + j] + H[2 * 3 + i] * H[2 * 3 + j];
    double scale = A[0][0] + A[1][1] + A[2][2];
    if (!(scale > 0.0)) {
        for (int i = 0; i < 9; ++i) R[i] = (i % 4 == 0) ? 1.0 : 0.0;
        return 2;
    }
    for (int sweep = 0; sweep < 12; ++sweep) {
        double off = fabs(A[0][1]) + fabs(A[0][2]) + fabs(A[1][2]);
        if (off <= 1e-30 * scale) break;
#pragma unroll
        for (int pq = 0; pq < 3; ++pq) {
            const int p = (pq == 2) ? 1 : 0;
            const int q = (pq == 0) ? 1 : 2;
            double apq = A[p][q];
            if (fabs(apq) <= 1e-300) continue;
            double theta = (A[q][q] - A[p][p]) / (2.0 * apq);
            double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
            double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
            // A <- J^T A J
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                double arp = A[r][p], arq = A[r][q];
                A[r][p] = c * arp - s * arq;
                A[r][q] = s * arp + c * arq;
            }
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                double apr = A[p][r], aqr = A[q][r];
                A[p][r] = c * apr - s * aqr;
                A[q][r] = s * apr + c * aqr;
            }
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                double vrp = V[r][p], vrq = V[r][q];
                V[r][p] = c * vrp - s * vrq;
                V[r][q] = s * vrp + c * vrq;
            }
        }
    }
    // order the two largest eigenvalues first
    int i0 = 0, i1 = 1, i2 = 2;
    double e0 = A[0][0], e1 = A[1][1], e2 = A[2][2];
    if (e0 < e1) { double t = e0; e0 = e1; e1 = t; int ti = i0; i0 = i1; i1 = ti; }
    if (e0 < e2) { double t = e0; e0 = e2; e2 = t; int ti = i0; i0 = i2; i2 = ti; }
    if (e1 < e2) { double t = e1; e1 = e2; e2 = t; int ti = i1; i1 = i2; i2 = ti; }
    (void)i2;
    double v1[3] = {V[0][i0], V[1][i0], V[2][i0]};
    double v2[3] = {V[0][i1], V[1][i1], V[2][i1]};
    double v3[3] = {v1[1] * v2[2] - v1[2] * v2[1], v1[2] * v2[0] - v1[0] * v2[2], v1[0] * v2[1] - v1[1] * v2[0]};
    double u1[3], u2[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        u1[r] = H[r * 3 + 0] * v1[0] + H[r * 3 + 1] * v1[1] + H[r * 3 + 2] * v1[2];
        u2[r] = H[r * 3 + 0] * v2[0] + H[r * 3 + 1] * v2[1] + H[r * 3 + 2] * v2[2];
    }
    int flag = 0;
    double n1 = sqrt(u1[0] * u1[0] + u1[1] * u1[1] + u1[2] * u1[2]);
    u1[0] /= n1; u1[1] /= n1; u1[2] /= n1;
    double d = u1[0] * u2[0] + u1[1] * u2[1] + u1[2] * u2[2];
    u2[0] -= d * u1[0]; u2[1] -= d * u1[1]; u2[2] -= d * u1[2];
    double n2 = sqrt(u2[0] * u2[0] + u2[1] * u2[1] + u2[2] * u2[2]);
    if (n2 <= 1e-7 * n1) {
        // rank one: any unit vector orthogonal to u1 (the reference's choice is LAPACK's whim)
        flag = 1;
        int m = (fabs(u1[0]) <= fabs(u1[1]) && fabs(u1[0]) <= fabs(u1[2])) ? 0 : (fabs(u1[1]) <= fabs(u1[2]) ? 1 : 2);
        double e[3] = {0, 0, 0};
        e[m] = 1.0;
        double dd = u1[m];
        u2[0] = e[0] - dd * u1[0]; u2[1] = e[1] - dd * u1[1]; u2[2] = e[2] - dd * u1[2];
        n2 = sqrt(u2[0] * u2[0] + u2[1] * u2[1] + u2[2] * u2[2]);
    }
    u2[0] /= n2; u2[1] /= n2; u2[2] /= n2;
    double u3[3] = {u1[1] * u2[2] - u1[2] * u2[1], u1[2] * u2[0] - u1[0] * u2[2], u1[0] * u2[1] - u1[1] * u2[0]};
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) R[i * 3 + j] = v1[i] * u1[j] + v2[i] * u2[j] + v3[i] * u3[j];
    return flag;
}

// Assemble [R t; 0 1] with t = cB - R cA (models/common.py:42, utils/SE3.py:86-88).
__device__ inline void write_transform(float* T, const double R[9], const float cA[3], const float cB[3]) {
    float Rf[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) Rf[i] = (float)R[i];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        float rc = __fadd_rn(__fadd_rn(__fmul_rn(Rf[i * 3 + 0], cA[0]), __fmul_rn(Rf[i * 3 + 1], cA[1])),
                             __fmul_rn(Rf[i * 3 + 2], cA[2]));
        T[i * 4 + 0] = Rf[i * 3 + 0];
        T[i * 4 + 1] = Rf[i * 3 + 1];
        T[i * 4 + 2] = Rf[i * 3 + 2];
        T[i * 4 + 3] = cB[i] - rc;
    }
    T[12] = 0.f; T[13] = 0.f; T[14] = 0.f; T[15] = 1.f;
}

// Squared-distance threshold equivalent to comparing the fp32 norm with R: sqrtf is monotone and correctly rounded, so
// sqrtf(d2) >= R  <=>  d2 >= T2 (and sqrtf(d2) < R  <=>  d2 < T2) with T2 the smallest fp32 whose square root reaches R.
inline float squared_threshold(float R) {
    if (!(R > 0.f)) return 0.f;                       // every distance is >= R
    float T2 = R * R;
    while (T2 > 0.f && sqrtf(T2) >= R) T2 = nextafterf(T2, 0.f);
    while (sqrtf(T2) < R) T2 = nextafterf(T2, INFINITY);
    return T2;
}

// ---- launchers implemented in the .cu files -------------------------------------------------
int launch_score_hypotheses(const float* seed_trans, const float* src, const float* tgt, int B, int N, int S,
                            float thr, int32_t* inlier_count, int64_t* best_idx, float* final_trans,
                            float* final_labels, cudaStream_t st);
int launch_post_refine(float* trans, const float* src, const float* tgt, int B, int N, float tau, int max_iter,
                       int32_t* iters_out, int32_t* counts_out, cudaStream_t st);
int launch_feature_head(const float* feat, long long rows, int C, int H1, int H2, const float* w1t, const float* b1, const float* w2t,
                        const float* b2, const float* w3, const float* b3, float eps, float* normed, float* conf, cudaStream_t st);
size_t pick_seeds_workspace(int B, int N);
int launch_pick_seeds(const float* src, const float* scores, int B, int N, float R, int S, int64_t* seeds,
                      uint8_t* is_local_max, int32_t* n_pos, void* ws, size_t ws_bytes, cudaStream_t st);
size_t seed_knn_workspace(int B, int N, int S);
int launch_seed_knn(const float* normed, const int64_t* seeds, int B, int N, int C, int S, int k,
                    int64_t* knn_idx, float* knn_dist, void* ws, size_t ws_bytes, cudaStream_t st);
size_t seed_hypotheses_workspace(int B, int S, int k, int num_iterations);
int launch_seed_hypotheses(const float* normed, const float* src, const float* tgt, const int64_t* knn_idx, int B,
                           int N, int C, int S, int k, float sigma, float sigma_spat, int num_iterations,
                           float* seed_trans, float* weights, int32_t* iters, void* ws, size_t ws_bytes,
                           cudaStream_t st);
int launch_attn_fp32(const float* q, const float* k, const float* v, int64_t ld, const float* src, const float* tgt,
                     float* out, int64_t ld_out, int B, int N, int C, float sigma_spat, cudaStream_t st);
int launch_nn_match(const float* A, const float* Bm, int P, int Na, int Nb, int D, int metric, int64_t* idx1, int64_t* idx2,
                    float* val1, float* val2, cudaStream_t st);
int launch_mutual_corr(const int64_t* idx_ab, const int64_t* idx_ba, int P, int Na, int Nb, int use_mutual, int64_t* corr,
                       int32_t* count, cudaStream_t st);
int launch_registration_metrics(const float* trans, const float* gt, const float* src, const float* tgt, int P, int N, float* out,
                                cudaStream_t st);
int launch_classification_counts(const float* pred, const float* gt, int P, int N, float thr, int32_t* counts, float* sums,
                                 cudaStream_t st);
size_t attn_pp_workspace(int B, int N);
int launch_attn_pp(const float* q, const float* k, const float* v, int64_t ld, const float* src, const float* tgt,
                   float* out, int64_t ld_out, int B, int N, int C, float sigma_spat, int mode, void* ws,
                   size_t ws_bytes, cudaStream_t st);
int launch_attn_pp_main(float* out, int64_t ld_out, int B, int N, int C, int mode, void* ws, size_t ws_bytes, cudaStream_t st);
int launch_qkv_pack(const float* y, const void* wpack, const float* bias, const float* src, const float* tgt, int B, int N,
                    int C, float sigma_spat, void* ws, size_t ws_bytes, cudaStream_t st);
int launch_compat_matvec(const float* src, const float* tgt, const float* v, float* y, int N, int row_begin,
                         int row_end, float sigma_spat, const int32_t* converged_flag, cudaStream_t st);
int launch_split16_rows(const float* x, const float* add, float* y_out, void* out16, long long rows, int K, int Kp,
                        int relu, cudaStream_t st);
int launch_mlp_chain(const float* x, const float* add_in, long long rows, int n_layers, const int32_t* dims,
                     const int32_t* relu, const void* wpack, const float* bias, const float* add_out, float* y,
                     void* y_packed, int Kp, cudaStream_t st);
int launch_mlp_chain_direct(const float* x, long long rows, int n_layers, const int32_t* dims, const int32_t* relu, const void* wpack,
                            const float* bias, const float* res, int res_layer, float* y, cudaStream_t st);
size_t power_workspace(int N);
size_t power_p2p_exchange_bytes(int N);
int launch_compat_matvec_p2p(const float* src, const float* tgt, const float* v, int N, int row_begin, int row_end,
                             float sigma_spat, const int32_t* converged_flag, void* const* peer_bases, int world, int rank,
                             int seq, cudaStream_t st);
int launch_power_normalize_p2p(void* exchange_base, float* v, int N, int32_t* converged_flag, int32_t* iters_done, int world,
                               int seq, cudaStream_t st);
int power_iter_full_static(const float* src, const float* tgt, float* v_out, int N, float sigma_spat, int num_iterations,
                           int32_t* iters_out, cudaStream_t st);
int power_iter_full(const float* src, const float* tgt, float* v_out, int N, float sigma_spat, int num_iterations,
                    int32_t* iters_out, void* ws, size_t ws_bytes, cudaStream_t st);
int launch_power_normalize(const float* y, float* v, int N, int32_t* converged_flag, int32_t* iters_done, void* ws,
                           size_t ws_bytes, cudaStream_t st);

size_t frontend_search_workspace(int n);
size_t frontend_voxel_workspace(int n, int max_cells);
int launch_hybrid_search(const double* pts, int n, double radius, int max_nn, int32_t* idx, double* d2, int32_t* cnt,
                         double* normals, void* ws, size_t ws_bytes, cudaStream_t st);
int launch_voxel_down_sample(const double* pts, const double* normals, int n, double voxel, double* out_pts,
                             double* out_normals, int32_t* out_count, void* ws, size_t ws_bytes, cudaStream_t st);
int launch_fpfh(const double* pts, const double* normals, int n, const int32_t* idx, const double* d2, const int32_t* cnt,
                int max_nn, double* spfh, double* out, cudaStream_t st);
size_t icp_workspace(int B, int N);
int launch_icp_refine(const float* src, const float* tgt, float* trans, int B, int N, double max_dist, int max_iter,
                      double rel_fitness, double rel_rmse, double* stats, int32_t* iters, int32_t* corr, void* ws, size_t ws_bytes,
                      cudaStream_t st);
size_t nn_match_f64_workspace(int Na);
int launch_nn_match_f64(const double* A, const double* Bm, int Na, int Nb, int D, int64_t* idx, double* val, void* ws,
                        size_t ws_bytes, cudaStream_t st);

}  // namespace pdsc
