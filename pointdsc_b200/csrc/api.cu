// C ABI of libpointdsc_b200.so (include/pointdsc_b200.h): argument checks, error strings,
// workspace sizing.  No torch types cross this boundary.
#include "common.cuh"

#include <stdlib.h>

#include <atomic>
#include <stdarg.h>
#include <string.h>

namespace pdsc {

static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};

void count_launches(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        return PDSC_ERR_CUDA;
    }
    return PDSC_OK;
}

int num_sms() {
    static thread_local int cached_dev = -1, cached = 0;
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev != cached_dev) {
        cudaDeviceGetAttribute(&cached, cudaDevAttrMultiProcessorCount, dev);
        cached_dev = dev;
    }
    return cached > 0 ? cached : 148;
}

#ifdef PDSC_TUNING
int tuning_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return (e && e[0]) ? atoi(e) : dflt;
}
const char* tuning_str(const char* name) {
    const char* e = getenv(name);
    return (e && e[0]) ? e : nullptr;
}
#endif

static thread_local int32_t* g_overflow_flag = nullptr;
int32_t* overflow_flag() { return g_overflow_flag; }

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace pdsc

using namespace pdsc;

extern "C" {

int pdsc_version(void) { return 100; }

const char* pdsc_last_error(void) { return g_err; }

uint64_t pdsc_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int pdsc_check_device(void) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { set_error("no CUDA device"); return PDSC_ERR_DEVICE; }
    int major = 0, minor = 0;
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
    if (major != 10) {
        set_error("device %d is sm_%d%d; this library is built for sm_100a (B200) only", dev, major, minor);
        return PDSC_ERR_DEVICE;
    }
    return PDSC_OK;
}

size_t pdsc_workspace_bytes(int op, int B, int N, int C, int S, int k, int mode) {
    switch (op) {
        case PDSC_OP_NONLOCAL_ATTN: {
            if (mode == PDSC_ATTN_FP32) return 0;
            return attn_pp_workspace(B, N);
        }
        case PDSC_OP_PICK_SEEDS: return pick_seeds_workspace(B, N);
        case PDSC_OP_HYBRID_SEARCH: return frontend_search_workspace(N);
        case PDSC_OP_VOXEL_DOWN_SAMPLE: return frontend_voxel_workspace(N, S);
        case PDSC_OP_NN_MATCH_F64: return nn_match_f64_workspace(N);
        case PDSC_OP_ICP_REFINE: return icp_workspace(B, N);
        case PDSC_OP_SEED_KNN: return seed_knn_workspace(B, N, S);
        case PDSC_OP_SEED_HYPOTHESES: return seed_hypotheses_workspace(B, S, k, 31);
        case PDSC_OP_SCORE_HYPOTHESES: return 0;
        case PDSC_OP_POST_REFINE: return 0;
        case PDSC_OP_POWER_ITER_FULL: return power_workspace(N);
        default: return 0;
    }
}

int pdsc_nonlocal_attn(const float* q, const float* k, const float* v, int64_t ld_qkv, const float* src,
                       const float* tgt, float* out, int64_t ld_out, int B, int N, int C, float sigma_spat, int mode,
                       void* ws, size_t ws_bytes, void* stream) {
    PDSC_REQUIRE(q && k && v && src && tgt && out, "nonlocal_attn: null pointer");
    PDSC_REQUIRE(B > 0 && N > 0, "nonlocal_attn: bad shape B=%d N=%d", B, N);
    PDSC_REQUIRE(ld_qkv >= C && ld_out >= C && ld_qkv % 4 == 0 && ld_out % 4 == 0, "nonlocal_attn: bad row stride");
    PDSC_REQUIRE(aligned16(q) && aligned16(k) && aligned16(v) && aligned16(out), "nonlocal_attn: q/k/v/out must be 16-byte aligned");
    PDSC_REQUIRE(sigma_spat > 0.f, "nonlocal_attn: sigma_spat must be positive");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (mode == PDSC_ATTN_FP32) return launch_attn_fp32(q, k, v, ld_qkv, src, tgt, out, ld_out, B, N, C, sigma_spat, st);
    if (mode == PDSC_ATTN_SPLIT16 || mode == PDSC_ATTN_FP16 || mode == PDSC_ATTN_SPLIT16_PV2)
        return launch_attn_pp(q, k, v, ld_qkv, src, tgt, out, ld_out, B, N, C, sigma_spat, mode, ws, ws_bytes, st);
    set_error("nonlocal_attn: unknown precision mode %d", mode);
    return PDSC_ERR_ARG;
}

int pdsc_qkv_pack(const float* x, const void* wpack, const float* bias, const float* src, const float* tgt, int B, int N,
                  int C, float sigma_spat, void* ws, size_t ws_bytes, void* stream) {
    PDSC_REQUIRE(x && wpack && bias && src && tgt && ws, "qkv_pack: null pointer");
    PDSC_REQUIRE(B > 0 && N > 0, "qkv_pack: bad shape B=%d N=%d", B, N);
    PDSC_REQUIRE(aligned16(x) && aligned16(wpack), "qkv_pack: x and wpack must be 16-byte aligned");
    PDSC_REQUIRE(sigma_spat > 0.f, "qkv_pack: sigma_spat must be positive");
    return launch_qkv_pack(x, wpack, bias, src, tgt, B, N, C, sigma_spat, ws, ws_bytes, reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_nonlocal_attn_packed(float* out, int64_t ld_out, int B, int N, int C, int mode, void* ws, size_t ws_bytes,
                              void* stream) {
    PDSC_REQUIRE(out && ws, "nonlocal_attn_packed: null pointer");
    PDSC_REQUIRE(B > 0 && N > 0, "nonlocal_attn_packed: bad shape B=%d N=%d", B, N);
    PDSC_REQUIRE(ld_out >= C && ld_out % 4 == 0 && aligned16(out), "nonlocal_attn_packed: bad output stride/alignment");
    PDSC_REQUIRE(mode == PDSC_ATTN_SPLIT16 || mode == PDSC_ATTN_FP16 || mode == PDSC_ATTN_SPLIT16_PV2,
                 "nonlocal_attn_packed: mode must be SPLIT16, SPLIT16_PV2 or FP16");
    return launch_attn_pp_main(out, ld_out, B, N, C, mode, ws, ws_bytes, reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_nn_match(const float* A, const float* B, int P, int Na, int Nb, int D, int metric, int64_t* idx1, int64_t* idx2,
                  float* val1, float* val2, void* stream) {
    PDSC_REQUIRE(A && B && idx1, "nn_match: null pointer");
    PDSC_REQUIRE(P > 0 && Na > 0 && Nb > 0 && D > 0, "nn_match: bad shape P=%d Na=%d Nb=%d D=%d", P, Na, Nb, D);
    PDSC_REQUIRE(P <= 65535, "nn_match: at most 65535 problems per call");
    PDSC_REQUIRE(metric == PDSC_MATCH_DOT || metric == PDSC_MATCH_L2, "nn_match: unknown metric %d", metric);
    return launch_nn_match(A, B, P, Na, Nb, D, metric, idx1, idx2, val1, val2, reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_mutual_corr(const int64_t* idx_ab, const int64_t* idx_ba, int P, int Na, int Nb, int use_mutual, int64_t* corr,
                     int32_t* count, void* stream) {
    PDSC_REQUIRE(idx_ab && corr && count && (idx_ba || !use_mutual), "mutual_corr: null pointer");
    PDSC_REQUIRE(P > 0 && Na > 0 && Nb > 0, "mutual_corr: bad shape P=%d Na=%d Nb=%d", P, Na, Nb);
    return launch_mutual_corr(idx_ab, idx_ba, P, Na, Nb, use_mutual, corr, count, reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_registration_metrics(const float* trans, const float* gt_trans, const float* src, const float* tgt, int P, int N,
                              float* out, void* stream) {
    PDSC_REQUIRE(trans && gt_trans && src && tgt && out, "registration_metrics: null pointer");
    PDSC_REQUIRE(P > 0 && N > 0, "registration_metrics: bad shape P=%d N=%d", P, N);
    return launch_registration_metrics(trans, gt_trans, src, tgt, P, N, out, reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_classification_counts(const float* pred, const float* gt, int P, int N, float thr, int32_t* counts, float* sums,
                               void* stream) {
    PDSC_REQUIRE(pred && gt && counts && sums, "classification_counts: null pointer");
    PDSC_REQUIRE(P > 0 && N > 0, "classification_counts: bad shape P=%d N=%d", P, N);
    return launch_classification_counts(pred, gt, P, N, thr, counts, sums, reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_mlp_chain(const float* x, const float* add_in, int64_t rows, int n_layers, const int32_t* dims,
                   const int32_t* relu, const void* wpack, const float* bias, const float* add_out, float* y_out,
                   void* y_packed, int Kp, void* stream) {
    PDSC_REQUIRE(x && dims && relu && wpack && bias && (y_out || y_packed), "mlp_chain: null pointer");
    PDSC_REQUIRE(rows > 0, "mlp_chain: rows must be positive");
    PDSC_REQUIRE(aligned16(x) && aligned16(wpack) && (!add_in || aligned16(add_in)) && (!add_out || aligned16(add_out)) &&
                     (!y_out || aligned16(y_out)) && (!y_packed || aligned16(y_packed)),
                 "mlp_chain: pointers must be 16-byte aligned");
    return launch_mlp_chain(x, add_in, rows, n_layers, dims, relu, wpack, bias, add_out, y_out, y_packed, Kp,
                            reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_mlp_chain_direct(const float* x, int64_t rows, int n_layers, const int32_t* dims, const int32_t* relu, const void* wpack,
                          const float* bias, const float* residual, int residual_layer, float* y_out, void* stream) {
    PDSC_REQUIRE(x && dims && relu && wpack && bias && y_out, "mlp_chain_direct: null pointer");
    PDSC_REQUIRE(rows > 0, "mlp_chain_direct: rows must be positive");
    PDSC_REQUIRE(aligned16(x) && aligned16(wpack) && aligned16(y_out) && (!residual || aligned16(residual)),
                 "mlp_chain_direct: pointers must be 16-byte aligned");
    return launch_mlp_chain_direct(x, rows, n_layers, dims, relu, wpack, bias, residual, residual_layer, y_out,
                                   reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_pick_seeds(const float* src, const float* scores, int B, int N, float R, int S, int64_t* seeds_out,
                    uint8_t* is_local_max_out, int32_t* n_pos_out, void* ws, size_t ws_bytes, void* stream) {
    PDSC_REQUIRE(src && scores && seeds_out && ws, "pick_seeds: null pointer");
    PDSC_REQUIRE(B > 0 && N > 0 && S >= 0 && S <= N, "pick_seeds: bad shape B=%d N=%d S=%d", B, N, S);
    if (S == 0) return PDSC_OK;
    return launch_pick_seeds(src, scores, B, N, R, S, seeds_out, is_local_max_out, n_pos_out, ws, ws_bytes,
                             reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_seed_knn(const float* normed, const int64_t* seeds, int B, int N, int C, int S, int k, int64_t* knn_idx_out,
                  float* knn_dist_out, void* ws, size_t ws_bytes, void* stream) {
    PDSC_REQUIRE(normed && seeds && knn_idx_out && ws, "seed_knn: null pointer");
    PDSC_REQUIRE(B > 0 && N > 1 && S > 0, "seed_knn: bad shape B=%d N=%d S=%d", B, N, S);
    PDSC_REQUIRE(aligned16(normed) && C % 4 == 0, "seed_knn: features must be 16-byte aligned, C %% 4 == 0");
    return launch_seed_knn(normed, seeds, B, N, C, S, k, knn_idx_out, knn_dist_out, ws, ws_bytes,
                           reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_seed_hypotheses(const float* normed, const float* src, const float* tgt, const int64_t* knn_idx, int B, int N,
                         int C, int S, int k, float sigma, float sigma_spat, int num_iterations, float* seed_trans_out,
                         float* weights_out, int32_t* iters_out, void* ws, size_t ws_bytes, void* stream) {
    PDSC_REQUIRE(normed && src && tgt && knn_idx && seed_trans_out && ws, "seed_hypotheses: null pointer");
    PDSC_REQUIRE(B > 0 && N > 1 && S > 0, "seed_hypotheses: bad shape B=%d N=%d S=%d", B, N, S);
    PDSC_REQUIRE(aligned16(normed) && aligned16(seed_trans_out), "seed_hypotheses: features/transforms must be 16-byte aligned");
    return launch_seed_hypotheses(normed, src, tgt, knn_idx, B, N, C, S, k, sigma, sigma_spat, num_iterations,
                                  seed_trans_out, weights_out, iters_out, ws, ws_bytes,
                                  reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_score_hypotheses(const float* seed_trans, const float* src, const float* tgt, int B, int N, int S, float thr,
                          int32_t* inlier_count_out, int64_t* best_idx_out, float* final_trans_out,
                          float* final_labels_out, void* ws, size_t ws_bytes, void* stream) {
    (void)ws; (void)ws_bytes;
    PDSC_REQUIRE(seed_trans && src && tgt && inlier_count_out && best_idx_out && final_trans_out && final_labels_out,
                 "score_hypotheses: null pointer");
    PDSC_REQUIRE(B > 0 && N > 0 && S > 0, "score_hypotheses: bad shape B=%d N=%d S=%d", B, N, S);
    PDSC_REQUIRE(aligned16(seed_trans), "score_hypotheses: seed_trans must be 16-byte aligned");
    return launch_score_hypotheses(seed_trans, src, tgt, B, N, S, thr, inlier_count_out, best_idx_out, final_trans_out,
                                   final_labels_out, reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_post_refine(float* trans_inout, const float* src, const float* tgt, int B, int N, float tau, int max_iter,
                     int32_t* iters_out, int32_t* counts_out, void* stream) {
    PDSC_REQUIRE(trans_inout && src && tgt, "post_refine: null pointer");
    PDSC_REQUIRE(B > 0 && N > 0 && max_iter >= 0 && tau > 0.f, "post_refine: bad arguments");
    return launch_post_refine(trans_inout, src, tgt, B, N, tau, max_iter, iters_out, counts_out,
                              reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_feature_head(const float* feat, int64_t rows, int C, int H1, int H2, const float* w1t, const float* b1, const float* w2t,
                      const float* b2, const float* w3, const float* b3, float eps, float* normed_out, float* conf_out, void* stream) {
    PDSC_REQUIRE(feat && w1t && b1 && w2t && b2 && w3 && b3 && normed_out && conf_out, "feature_head: null pointer");
    PDSC_REQUIRE(rows > 0, "feature_head: bad row count");
    PDSC_REQUIRE(aligned16(feat) && aligned16(normed_out) && aligned16(w1t) && aligned16(w2t), "feature_head: pointers must be 16-byte aligned");
    return launch_feature_head(feat, rows, C, H1, H2, w1t, b1, w2t, b2, w3, b3, eps, normed_out, conf_out,
                               reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_split16_rows(const float* x, const float* add, float* y_out, void* out16, int64_t rows, int K, int Kp, int relu,
                      void* stream) {
    PDSC_REQUIRE(x && out16, "split16_rows: null pointer");
    PDSC_REQUIRE(rows > 0 && K > 0 && K % 8 == 0 && Kp % 8 == 0 && Kp >= 3 * K + 2, "split16_rows: need K %% 8 == 0, Kp %% 8 == 0, Kp >= 3K+2");
    PDSC_REQUIRE(aligned16(x) && aligned16(out16) && (!add || aligned16(add)) && (!y_out || aligned16(y_out)), "split16_rows: pointers must be 16-byte aligned");
    return launch_split16_rows(x, add, y_out, out16, rows, K, Kp, relu, reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_compat_matvec(const float* src, const float* tgt, const float* v, float* y, int N, int row_begin, int row_end,
                       float sigma_spat, const int32_t* converged_flag, void* stream) {
    PDSC_REQUIRE(src && tgt && v && y, "compat_matvec: null pointer");
    PDSC_REQUIRE(N > 0 && row_begin >= 0 && row_end <= N && row_begin <= row_end, "compat_matvec: bad row range");
    return launch_compat_matvec(src, tgt, v, y, N, row_begin, row_end, sigma_spat, converged_flag,
                                reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_power_normalize(const float* y, float* v_inout, int N, int32_t* converged_flag, int32_t* iters_done, void* ws,
                         size_t ws_bytes, void* stream) {
    PDSC_REQUIRE(y && v_inout && converged_flag, "power_normalize: null pointer");
    return launch_power_normalize(y, v_inout, N, converged_flag, iters_done, ws, ws_bytes,
                                  reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_power_iter_full(const float* src, const float* tgt, float* v_out, int N, float sigma_spat, int num_iterations,
                         int32_t* iters_out, void* ws, size_t ws_bytes, void* stream) {
    PDSC_REQUIRE(src && tgt && v_out && ws, "power_iter_full: null pointer");
    PDSC_REQUIRE(N > 0 && num_iterations > 0, "power_iter_full: bad arguments");
    return power_iter_full(src, tgt, v_out, N, sigma_spat, num_iterations, iters_out, ws, ws_bytes,
                           reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_set_overflow_flag(int32_t* device_flag) {
    g_overflow_flag = device_flag;
    return PDSC_OK;
}

size_t pdsc_power_p2p_exchange_bytes(int N) { return N > 0 ? power_p2p_exchange_bytes(N) : 0; }

int pdsc_compat_matvec_p2p(const float* src, const float* tgt, const float* v, int N, int row_begin, int row_end,
                           float sigma_spat, const int32_t* converged_flag, void* const* peer_bases, int world, int rank, int seq,
                           void* stream) {
    PDSC_REQUIRE(src && tgt && v && peer_bases, "compat_matvec_p2p: null pointer");
    PDSC_REQUIRE(N > 0 && row_begin >= 0 && row_end <= N && row_begin <= row_end, "compat_matvec_p2p: bad row range");
    PDSC_REQUIRE(world >= 1 && world <= PDSC_P2P_MAX_WORLD && rank >= 0 && rank < world && seq > 0,
                 "compat_matvec_p2p: bad world=%d rank=%d seq=%d (world <= %d, seq > 0)", world, rank, seq, PDSC_P2P_MAX_WORLD);
    for (int p = 0; p < world; ++p) PDSC_REQUIRE(peer_bases[p] && aligned16(peer_bases[p]), "compat_matvec_p2p: bad peer base %d", p);
    return launch_compat_matvec_p2p(src, tgt, v, N, row_begin, row_end, sigma_spat, converged_flag, peer_bases, world, rank, seq,
                                    reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_power_normalize_p2p(void* exchange_base, float* v_inout, int N, int32_t* converged_flag, int32_t* iters_done, int world,
                             int seq, void* stream) {
    PDSC_REQUIRE(exchange_base && v_inout && converged_flag, "power_normalize_p2p: null pointer");
    PDSC_REQUIRE(N > 0 && world >= 1 && world <= PDSC_P2P_MAX_WORLD && seq > 0, "power_normalize_p2p: bad arguments");
    return launch_power_normalize_p2p(exchange_base, v_inout, N, converged_flag, iters_done, world, seq,
                                      reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_power_iter_full_static(const float* src, const float* tgt, float* v_out, int N, float sigma_spat, int num_iterations,
                                int32_t* iters_out, void* stream) {
    PDSC_REQUIRE(src && tgt && v_out, "power_iter_full_static: null pointer");
    PDSC_REQUIRE(N > 0 && num_iterations > 0, "power_iter_full_static: bad arguments");
    return power_iter_full_static(src, tgt, v_out, N, sigma_spat, num_iterations, iters_out, reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_hybrid_search(const double* pts, int n, double radius, int max_nn, int32_t* idx_out, double* dist2_out,
                       int32_t* count_out, double* normals_out, void* ws, size_t ws_bytes, void* stream) {
    PDSC_REQUIRE(pts && ws && (idx_out || normals_out), "hybrid_search: null pointer");
    PDSC_REQUIRE(n > 0 && max_nn > 0 && max_nn <= 128 && radius > 0.0, "hybrid_search: bad arguments n=%d max_nn=%d", n, max_nn);
    PDSC_REQUIRE(!(dist2_out && !idx_out), "hybrid_search: dist2_out needs idx_out");
    return launch_hybrid_search(pts, n, radius, max_nn, idx_out, dist2_out, count_out, normals_out, ws, ws_bytes,
                                reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_voxel_down_sample(const double* pts, const double* normals, int n, double voxel, double* out_pts,
                           double* out_normals, int32_t* out_count, void* ws, size_t ws_bytes, void* stream) {
    PDSC_REQUIRE(pts && out_pts && out_count && ws, "voxel_down_sample: null pointer");
    PDSC_REQUIRE(!normals || out_normals, "voxel_down_sample: normals given but out_normals is null");
    PDSC_REQUIRE(n > 0 && voxel > 0.0, "voxel_down_sample: bad arguments n=%d", n);
    return launch_voxel_down_sample(pts, normals, n, voxel, out_pts, out_normals, out_count, ws, ws_bytes,
                                    reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_fpfh(const double* pts, const double* normals, int n, const int32_t* idx, const double* dist2,
              const int32_t* count, int max_nn, double* spfh_ws, double* fpfh_out, void* stream) {
    PDSC_REQUIRE(pts && normals && idx && dist2 && count && spfh_ws && fpfh_out, "fpfh: null pointer");
    PDSC_REQUIRE(n > 0 && max_nn > 0, "fpfh: bad arguments n=%d max_nn=%d", n, max_nn);
    return launch_fpfh(pts, normals, n, idx, dist2, count, max_nn, spfh_ws, fpfh_out, reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_nn_match_f64(const double* A, const double* B, int Na, int Nb, int D, int64_t* idx_out, double* val_out,
                      void* ws, size_t ws_bytes, void* stream) {
    PDSC_REQUIRE(A && B && idx_out && ws, "nn_match_f64: null pointer");
    PDSC_REQUIRE(Na > 0 && Nb > 0 && D > 0 && D <= 64, "nn_match_f64: bad shape Na=%d Nb=%d D=%d", Na, Nb, D);
    return launch_nn_match_f64(A, B, Na, Nb, D, idx_out, val_out, ws, ws_bytes, reinterpret_cast<cudaStream_t>(stream));
}

int pdsc_icp_refine(const float* src, const float* tgt, float* trans_inout, int B, int N, double max_dist, int max_iter,
                    double rel_fitness, double rel_rmse, double* stats_out, int32_t* iters_out, int32_t* corr_out,
                    void* ws, size_t ws_bytes, void* stream) {
    PDSC_REQUIRE(src && tgt && trans_inout && ws, "icp_refine: null pointer");
    PDSC_REQUIRE(B > 0 && N > 0 && max_dist > 0.0 && max_iter >= 0 && max_iter <= 1000, "icp_refine: bad arguments B=%d N=%d", B, N);
    return launch_icp_refine(src, tgt, trans_inout, B, N, max_dist, max_iter, rel_fitness, rel_rmse, stats_out, iters_out,
                             corr_out, ws, ws_bytes, reinterpret_cast<cudaStream_t>(stream));
}

}  // extern "C"
