/*
 * pointdsc_b200 -- C ABI of the B200 (sm_100a) kernels behind PointDSC.forward(data).
 *
 * The reference (AmnonDrory/PointDSC) is pure Python/PyTorch and has no FFI: the interface
 * this library replaces is the set of Python methods listed beside each entry point
 * (file:line relative to the reference checkout).  A maintainer binds these symbols with
 * ctypes (INTEGRATION.md shows the stub); pointdsc_b200/_cabi.py is that binding.
 *
 * Conventions
 *  - Every pointer is a DEVICE pointer owned by the caller (inputs, outputs and workspace).
 *    Kernels never allocate or free and keep no pointer after return.
 *  - Tensors are dense row-major: keypoints [B,N,3] fp32, features [B,N,C] fp32 (channel last),
 *    transforms [B,S,4,4] fp32, indices int64 (what torch produces), counts int32.
 *  - Work is enqueued on `stream` (a cudaStream_t passed as void*), asynchronously; no
 *    device-wide synchronisation inside.  Reentrant across host threads/streams/devices.
 *  - Return value: PDSC_OK or an error code; pdsc_last_error() gives the message for the
 *    calling thread.  There is no CPU fallback anywhere.
 *  - B independent registration pairs per call (the reference's test path is batch-size 1,
 *    models/PointDSC.py:210,414; B>1 means B independent runs of it).
 */
#ifndef POINTDSC_B200_H
#define POINTDSC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PDSC_OK 0
#define PDSC_ERR_ARG 1         /* bad shape / null pointer / alignment */
#define PDSC_ERR_DEVICE 2      /* not an sm_100 device */
#define PDSC_ERR_CUDA 3        /* launch or runtime error, see pdsc_last_error */
#define PDSC_ERR_WORKSPACE 4   /* workspace too small */

/* precision modes of pdsc_nonlocal_attn */
#define PDSC_ATTN_FP32 0       /* CUDA-core fp32 FMA contractions (exact reference arithmetic) */
#define PDSC_ATTN_SPLIT16 1    /* tcgen05 fp16 hi/lo split, 3 MMAs per contraction (hi*hi + hi*lo + lo*hi), fp32 accumulate in TMEM */
#define PDSC_ATTN_FP16 2       /* tcgen05 single-pass fp16 (fast, outside the parity tolerance) */
#define PDSC_ATTN_SPLIT16_PV2 3 /* as SPLIT16 without the P_lo*V_hi term (row sum over the rounded P_hi): 6 % faster, message
                                  error ~2^-12 relative, i.e. OUTSIDE the parity tolerance on the KITTI-shaped batch (3e-4 in the
                                  confidences, profiles/r02_pv_terms.txt); kept for measurements, never the default */

/* operations for pdsc_workspace_bytes */
/* distance form of pdsc_nn_match */
#define PDSC_MATCH_DOT 0       /* sqrt(2 - 2 a.b + 1e-6)       (datasets/LidarFeatureExtractor.py:110) */
#define PDSC_MATCH_L2 1        /* sqrt(max(|a|^2+|b|^2-2 a.b, 1e-30))  (algorithms/matching.py:29-30)  */

#define PDSC_OP_NONLOCAL_ATTN 1
#define PDSC_OP_PICK_SEEDS 2
#define PDSC_OP_SEED_KNN 3
#define PDSC_OP_SEED_HYPOTHESES 4
#define PDSC_OP_SCORE_HYPOTHESES 5
#define PDSC_OP_POST_REFINE 6
#define PDSC_OP_POWER_ITER_FULL 7
#define PDSC_OP_HYBRID_SEARCH 8      /* N = number of points */
#define PDSC_OP_VOXEL_DOWN_SAMPLE 9  /* N = number of points, S = voxel-grid cells the workspace should hold */
#define PDSC_OP_NN_MATCH_F64 10      /* N = Na */
#define PDSC_OP_ICP_REFINE 11

int pdsc_version(void);
const char* pdsc_last_error(void);
/* Number of CUDA kernels this library has enqueued in this process (all threads). */
uint64_t pdsc_launch_count(void);
/* 0 when the current device is sm_100 (B200); PDSC_ERR_DEVICE otherwise */
int pdsc_check_device(void);
/* Bytes of caller-provided workspace an operation needs (256-byte aligned base expected).
 * `mode` is the attention precision mode for PDSC_OP_NONLOCAL_ATTN, else ignored. */
size_t pdsc_workspace_bytes(int op, int B, int N, int C, int S, int k, int mode);

/* (a) compatibility-weighted non-local attention; replaces
 *     models/PointDSC.py:150-153 (src_dist / corr_compatibility, never materialised) and
 *     NonLocalBlock.forward models/PointDSC.py:39-42 (QK^T/sqrt(C), softmax(compat*logits), PV).
 *  q,k,v: [B,N,C] fp32 with row stride ld_qkv floats (they may be slices of one [B,N,3C]
 *  projection output); src,tgt: [B,N,3]; out: message [B,N,C] fp32, row stride ld_out.
 *  sigma_spat is the fp32 parameter value; compat = max(0, 1 - (|si-sj|-|ti-tj|)^2 / sigma_spat^2). */
int pdsc_nonlocal_attn(const float* q, const float* k, const float* v, int64_t ld_qkv,
                       const float* src, const float* tgt, float* out, int64_t ld_out,
                       int B, int N, int C, float sigma_spat, int mode,
                       void* ws, size_t ws_bytes, void* stream);

/* Fused q/k/v projection + operand packing, and the attention on operands packed that way.
 *  pdsc_qkv_pack replaces projection_q/k/v of NonLocalBlock.forward (models/PointDSC.py:29-37): from the layer input
 *  x [B,N,C] fp32 it computes q = x Wq + bq, k, v on the tensor cores (fp16 hi/lo split, fp32 accumulation) and writes
 *  them into `ws` directly as the fp16 hi/lo operand tiles of the attention kernel (V already transposed), together
 *  with the sigma-normalised key points.  wpack (device): the tcgen05 operand images of Wq, Wk, Wv (layout as for
 *  pdsc_mlp_chain, 64 KB each); bias (device): bq|bk|bv fp32 [3C].  C must be 128.
 *  pdsc_nonlocal_attn_packed then produces the message [B,N,C] from `ws` (same ws, B, N, C; workspace size from
 *  pdsc_workspace_bytes(PDSC_OP_NONLOCAL_ATTN, ...)); mode PDSC_ATTN_SPLIT16 or PDSC_ATTN_FP16. */
int pdsc_qkv_pack(const float* x, const void* wpack, const float* bias, const float* src, const float* tgt,
                  int B, int N, int C, float sigma_spat, void* ws, size_t ws_bytes, void* stream);
int pdsc_nonlocal_attn_packed(float* out, int64_t ld_out, int B, int N, int C, int mode,
                              void* ws, size_t ws_bytes, void* stream);

/* Operand packing for the channel-last 1x1-conv GEMMs of NonLocalNet / NonLocalBlock
 *  (models/PointDSC.py:12-23,54-62; the GEMMs themselves stay torch/cuBLAS calls).
 *  v = x (+ add) (then ReLU if relu != 0), x/add/y_out fp32 [rows,K]; y_out (nullable) receives v;
 *  out16 fp16 [rows,Kp] receives [hi(v) | hi(v) | lo(v) | 1 1 0...] with hi+lo the two-fp16 split
 *  of v, so that out16 @ [W_hi; W_lo; W_hi; b_hi; b_lo; 0] (fp16 inputs, fp32 accumulate) = v W + b
 *  at fp32-grade accuracy.  K % 8 == 0, Kp % 8 == 0, Kp >= 3K+2. */
int pdsc_split16_rows(const float* x, const float* add, float* y_out, void* out16, int64_t rows, int K, int Kp,
                      int relu, void* stream);

/* Fused pointwise MLP chain for the channel-last 1x1-conv stacks (PointCN models/PointDSC.py:54-62,
 *  NonLocalBlock.fc_message + residual :12-18,44): y = L_{n-1}(..L_0(x + add_in)..) + add_out with
 *  L_i(v) = act_i(v W_i + b_i), computed on the tensor cores with fp16 hi/lo split operands and fp32 accumulation
 *  (fp32-grade accuracy), one 128-row tile kept on chip through the whole chain.
 *  x, add_in (nullable): fp32 [rows, dims[0]]; add_out (nullable), y_out (nullable): fp32 [rows, dims[n_layers]];
 *  y_packed (nullable): fp16 [rows, Kp] in the pdsc_split16_rows row layout (operand of a following library GEMM).
 *  dims (n_layers + 1 entries, each 64 or 128) and relu (n_layers flags) are HOST arrays; n_layers <= 3 (shared-memory staging of the rows; pdsc_mlp_chain_direct takes 4).
 *  wpack (device): per layer the tcgen05 B-operand image of W_i (BatchNorm folded): fp16 hi then fp16 lo, each
 *  dims[i]/64 blocks of [dims[i+1] rows x 64 input channels], 128-byte rows with the 128B swizzle (16-byte chunk c of
 *  row r stored at chunk c ^ (r & 7)).  bias (device): the b_i concatenated, fp32. */
int pdsc_mlp_chain(const float* x, const float* add_in, int64_t rows, int n_layers, const int32_t* dims,
                   const int32_t* relu, const void* wpack, const float* bias, const float* add_out,
                   float* y_out, void* y_packed, int Kp, void* stream);

/* The same chain with up to FOUR layers and the residual added after any layer:
 *      y = L_{n-1}( .. L_{r+1}( L_r( .. L_0(x) .. ) + residual ) .. ),   r = residual_layer (ignored when residual is NULL),
 *  which is fc_message + residual of one encoder layer followed by the PointCN of the next (NonLocalBlock.forward
 *  models/PointDSC.py:12-18,44, then PointCN :54-62) in one launch: the sum feat + fc_message(msg) is consumed on chip
 *  and never written.  Same arithmetic, term by term, as two pdsc_mlp_chain calls (results are bit-identical).
 *  x: fp32 [rows, dims[0]]; residual: fp32 [rows, dims[residual_layer + 1]]; y_out: fp32 [rows, dims[n_layers]];
 *  dims, relu, wpack, bias as for pdsc_mlp_chain; n_layers <= 4.  Rows move straight between global memory and registers
 *  (shared memory holds the weights only). */
int pdsc_mlp_chain_direct(const float* x, int64_t rows, int n_layers, const int32_t* dims, const int32_t* relu,
                          const void* wpack, const float* bias, const float* residual, int residual_layer,
                          float* y_out, void* stream);

/* (c) NMS seed picking; replaces PointDSC.pick_seeds models/PointDSC.py:199-217.
 *  is_local_max[i] = all_j (score_i >= score_j  or  |s_i-s_j| >= R);
 *  seeds = first S indices ordered by (score*is_local_max descending, index ascending).
 *  The reference's order among equal keys is unspecified (torch.argsort); ours is by index.
 *  seeds_out int64 [B,S]; is_local_max_out uint8 [B,N] (nullable); n_pos_out int32 [B]
 *  (nullable) = number of strictly positive keys (the prefix whose order is well defined). */
int pdsc_pick_seeds(const float* src, const float* scores, int B, int N, float R, int S,
                    int64_t* seeds_out, uint8_t* is_local_max_out, int32_t* n_pos_out,
                    void* ws, size_t ws_bytes, void* stream);

/* (c) feature-space kNN of the seed rows; replaces knn() models/common.py:48-69 as called at
 *  models/PointDSC.py:251-252 (ignore_self, normalized features) but only for the S seed rows.
 *  dist = 2 - 2 f_s.f_j; the k nearest j != s ordered by (dist ascending, index ascending).
 *  knn_idx_out int64 [B,S,k]; knn_dist_out fp32 [B,S,k] (nullable). */
int pdsc_seed_knn(const float* normed, const int64_t* seeds, int B, int N, int C, int S, int k,
                  int64_t* knn_idx_out, float* knn_dist_out,
                  void* ws, size_t ws_bytes, void* stream);

/* (d) per-seed local spectral matching + weighted Kabsch; replaces
 *  cal_seed_trans models/PointDSC.py:257-320, cal_leading_eigenvector :338-358 and
 *  rigid_transform_3d models/common.py:7-45 (3x3 SVD on device, no host round trip).
 *  sigma = learned feature bandwidth (self.sigma), sigma_spat = self.sigma_spat.
 *  The power iteration keeps the reference's stopping rule: all seeds of a pair stop at the
 *  first iteration where torch.allclose(v, v_last) holds over the whole [S,k] tensor.
 *  seed_trans_out fp32 [B,S,4,4]; weights_out fp32 [B,S,k] (nullable, normalised weights);
 *  iters_out int32 [B] (nullable, iterations used). */
int pdsc_seed_hypotheses(const float* normed, const float* src, const float* tgt,
                         const int64_t* knn_idx, int B, int N, int C, int S, int k,
                         float sigma, float sigma_spat, int num_iterations,
                         float* seed_trans_out, float* weights_out, int32_t* iters_out,
                         void* ws, size_t ws_bytes, void* stream);

/* (e) hypothesis scoring; replaces models/PointDSC.py:325-335.
 *  inlier_count[s] = #{n : |R_s src_n + t_s - tgt_n| < thr}; best = first argmax;
 *  final_trans = seed_trans[best]; final_labels = 0/1 fp32 mask under the best transform.
 *  inlier_count_out int32 [B,S]; best_idx_out int64 [B]; final_trans_out fp32 [B,4,4];
 *  final_labels_out fp32 [B,N]. */
int pdsc_score_hypotheses(const float* seed_trans, const float* src, const float* tgt,
                          int B, int N, int S, float thr,
                          int32_t* inlier_count_out, int64_t* best_idx_out,
                          float* final_trans_out, float* final_labels_out,
                          void* ws, size_t ws_bytes, void* stream);

/* (e) post refinement; replaces PointDSC.post_refinement models/PointDSC.py:403-438
 *  (+ transform utils/SE3.py:43-57, rigid_transform_3d models/common.py:7-45).
 *  Up to max_iter rounds of: d=|T src - tgt|, mask=d<tau, stop when the inlier count repeats,
 *  else refit T on the masked points with weights 1/(1+(d/tau)^2).  The whole loop runs on the
 *  device.  trans_inout fp32 [B,4,4]; iters_out int32 [B] (nullable; number of refits);
 *  counts_out int32 [B,max_iter] (nullable; inlier count seen at each round, -1 = not run). */
int pdsc_post_refine(float* trans_inout, const float* src, const float* tgt, int B, int N,
                     float tau, int max_iter, int32_t* iters_out, int32_t* counts_out,
                     void* stream);

/* Feature head: the two consumers of corr_features right after the encoder, in one pass over the rows --
 *  normed = F.normalize(corr_features, p=2, dim=1) (models/PointDSC.py:156; x / max(||x||_2, eps), eps = 1e-12) and
 *  confidence = self.classification(corr_features) (models/PointDSC.py:107-113 Conv1d C->H1, ReLU, H1->H2, ReLU, H2->1,
 *  called at :171).  feat fp32 [rows, C] channel-last; w1t [C,H1], w2t [H1,H2] (y = x @ wt), w3 [H2], b3 [1];
 *  normed_out fp32 [rows, C], conf_out fp32 [rows].  Built for C=128, H1=H2=32 (the reference's sizes); other
 *  sizes return PDSC_ERR_ARG (the Python host keeps torch ops for them). */
int pdsc_feature_head(const float* feat, int64_t rows, int C, int H1, int H2,
                      const float* w1t, const float* b1, const float* w2t, const float* b2,
                      const float* w3, const float* b3, float eps,
                      float* normed_out, float* conf_out, void* stream);

/* (b) full-matrix power iteration building block; cal_leading_eigenvector
 *  models/PointDSC.py:338-358 applied to the N x N compatibility matrix with zero diagonal
 *  (the call kept at models/PointDSC.py:170; baseline_scripts/baseline_3DMatch.py:19-46).
 *  The matrix is recomputed on the fly, never stored.
 *  pdsc_compat_matvec: y[r-row_begin] = sum_{j!=r} compat(r,j) v[j] for row_begin<=r<row_end
 *  (row sharding for multi-GPU: every rank owns a row range, then all-gathers y).
 *  pdsc_power_normalize: v_new = y/(|y|+1e-6); sets *converged_flag (int32, device) when
 *  allclose(v_new, v_last, rtol 1e-5, atol 1e-8); when the flag was already set on entry it
 *  leaves v unchanged (device-side early exit, no host sync).  iters_done is incremented
 *  on the device for every executed iteration.
 *  pdsc_power_iter_full: single-GPU driver, num_iterations launches of the pair above. */
int pdsc_compat_matvec(const float* src, const float* tgt, const float* v, float* y,
                       int N, int row_begin, int row_end, float sigma_spat,
                       const int32_t* converged_flag, void* stream);
int pdsc_power_normalize(const float* y, float* v_inout, int N, int32_t* converged_flag,
                         int32_t* iters_done, void* ws, size_t ws_bytes, void* stream);
int pdsc_power_iter_full(const float* src, const float* tgt, float* v_out, int N,
                         float sigma_spat, int num_iterations, int32_t* iters_out,
                         void* ws, size_t ws_bytes, void* stream);

/* Operand range of the tensor-core paths (pdsc_mlp_chain, pdsc_qkv_pack, pdsc_nonlocal_attn in the SPLIT16 / FP16 modes):
 *  every fp32 operand is split into fp16 hi + lo, so |value| must stay below 65504 (and relative accuracy degrades below
 *  ~0.1, absolute error floor ~3e-8).  There is no counterpart in the reference (it computes in fp32).  A kernel that meets
 *  a value outside the range ORs 1 into *device_flag (int32 in device memory owned by the caller, registered here for the
 *  calling host thread; NULL unregisters).  pdsc_feature_head raises the same flag for a confidence that is not finite.
 *  The flag is never cleared by the library. */
int pdsc_set_overflow_flag(int32_t* device_flag);

/* Row-sharded run of the same iteration over NVLink (BASELINE config 4: one pair, N >= 50k, rows split over the GPUs of
 *  one box; cal_leading_eigenvector models/PointDSC.py:338-358).  Instead of an all-gather per iteration, the matvec kernel
 *  of every rank stores its rows straight into the y buffer of EVERY rank (peer pointers into symmetric memory) and then
 *  publishes a system-scope arrival flag; the normalise kernel of every rank waits on the device for all `world` flags.
 *  Two launches per iteration, no collective, no host synchronisation; the bits equal the single-GPU run.
 *  Exchange buffer (one per rank, pdsc_power_p2p_exchange_bytes(N) bytes, zero-filled once before the first use, mapped
 *  on every peer): y[parity 0] | y[parity 1] | flags.  peer_bases: HOST array of `world` device pointers, entry p = the
 *  address of rank p's exchange buffer as seen from this rank.  seq: 1, 2, 3, ... one value per iteration, the same on
 *  all ranks, strictly increasing over the life of the buffers (the y parity is seq & 1).  All ranks must enqueue the
 *  same sequence of calls. */
#define PDSC_P2P_MAX_WORLD 8
size_t pdsc_power_p2p_exchange_bytes(int N);
int pdsc_compat_matvec_p2p(const float* src, const float* tgt, const float* v, int N, int row_begin, int row_end,
                           float sigma_spat, const int32_t* converged_flag, void* const* peer_bases,
                           int world, int rank, int seq, void* stream);
int pdsc_power_normalize_p2p(void* exchange_base, float* v_inout, int N, int32_t* converged_flag,
                             int32_t* iters_done, int world, int seq, void* stream);

/* No counterpart in the reference.  pdsc_power_iter_full for callers that
 *  cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): same kernels
 *  and same results, but the scratch (y[N], the converged flag, the iteration counter) is a file-scope device array
 *  inside the library instead of a caller-supplied workspace.  The entry point calls no cudaMalloc / cudaFree and no
 *  allocating library routine; the caller still owns src, tgt, v_out [N] and iters_out [1] (may be NULL).
 *  Limit: N <= PDSC_POWER_ITER_STATIC_MAX_N (larger N returns PDSC_ERR_ARG; use pdsc_power_iter_full).  Every GPU that
 *  loads the library holds the array (256 KB + 256 B) until exit.  All calls on a device share it and it keeps the last
 *  call's contents (each call resets what it needs): calls that overlap on the GPU (different streams) race; calls
 *  enqueued on one stream are safe. */
#define PDSC_POWER_ITER_STATIC_MAX_N 65536
int pdsc_power_iter_full_static(const float* src, const float* tgt, float* v_out, int N,
                                float sigma_spat, int num_iterations, int32_t* iters_out, void* stream);

/* ---- next rows: the callers either side of the forward -------------------------------------------- */

/* (N2) nearest neighbour in descriptor space; replaces find_nn algorithms/matching.py:22-65 and the
 *  argmin of datasets/LidarFeatureExtractor.py:110-111.  A fp32 [P,Na,D], B fp32 [P,Nb,D] (P independent problems,
 *  D <= 64).  idx1 int64 [P,Na]: index of the nearest B row (ties -> lowest index, like np.argmin / torch.min);
 *  idx2 (nullable): the second nearest (the reference masks the first and takes the min again, :35-37);
 *  val1 / val2 (nullable) fp32: the two distances, so a caller can see how close a decision was. */
int pdsc_nn_match(const float* A, const float* B, int P, int Na, int Nb, int D, int metric,
                  int64_t* idx1, int64_t* idx2, float* val1, float* val2, void* stream);

/* (N2) mutual check + compaction; replaces datasets/LidarFeatureExtractor.py:112-117 and nn_to_mutual /
 *  torch_intersect algorithms/matching.py:67-87,225-242.  idx_ab int64 [P,Na] (A -> B), idx_ba int64 [P,Nb]
 *  (B -> A; may be NULL when use_mutual == 0).  corr int64 [P,Na,2] receives the kept (i, idx_ab[i]) pairs in
 *  ascending i, count int32 [P] how many. */
int pdsc_mutual_corr(const int64_t* idx_ab, const int64_t* idx_ba, int P, int Na, int Nb, int use_mutual,
                     int64_t* corr, int32_t* count, void* stream);

/* (N3) registration metrics; replaces the per-pair loop of TransformationLoss.forward libs/loss.py:34-63.
 *  trans, gt_trans fp32 [P,4,4]; src, tgt fp32 [P,N,3].  out fp32 [P,4] = rotation error [deg], translation error
 *  [cm] (the reference's x100), mean |T src - tgt|, mean |T src - tgt|^2. */
int pdsc_registration_metrics(const float* trans, const float* gt_trans, const float* src, const float* tgt,
                              int P, int N, float* out, void* stream);

/* (N3) inlier classification statistics; replaces the sklearn calls of ClassificationLoss.forward
 *  libs/loss.py:96-112.  pred, gt fp32 [P,N]; predicted label = pred > thr, true label = gt != 0.
 *  counts int32 [P,4] = tp, fp, fn, tn; sums fp32 [P,2] = sum pred*gt, sum pred*(1-gt). */
int pdsc_classification_counts(const float* pred, const float* gt, int P, int N, float thr,
                               int32_t* counts, float* sums, void* stream);

/* ---- (N1) demo front end: what demo_registration.py:37-44,101-104 does through open3d 0.9 / numpy ----------------
 * float64 throughout, as open3d and the demo's numpy lines are.  One cloud per call. */

/* (N1) KDTreeSearchParamHybrid(radius, max_nn) for every point of the cloud against the cloud itself; replaces the
 *  KDTreeFlann searches inside estimate_normals (utils/pointcloud.py:20-21) and compute_fpfh_feature
 *  (demo_registration.py:41).  pts fp64 [n,3]; max_nn <= 128.  For query i: the up-to-max_nn points with
 *  dist^2 < radius^2 (i itself included), ascending (dist^2, index): idx_out int32 [n,max_nn] (-1 padded),
 *  dist2_out fp64 [n,max_nn] (+inf padded; nullable), count_out int32 [n] (nullable).
 *  normals_out (nullable) fp64 [n,3]: open3d's EstimateNormals on those neighbours -- covariance from the running
 *  sums, eigenvector of the smallest eigenvalue, not oriented; (0,0,1) with fewer than 3 neighbours.  With
 *  idx_out == NULL only the normals are produced and no neighbour list is written. */
int pdsc_hybrid_search(const double* pts, int n, double radius, int max_nn, int32_t* idx_out, double* dist2_out,
                       int32_t* count_out, double* normals_out, void* ws, size_t ws_bytes, void* stream);

/* (N1) PointCloud.voxel_down_sample(voxel) (demo_registration.py:40): voxel index floor((p - (min - voxel/2)) / voxel),
 *  mean point and normalised summed normal (nullable) per occupied voxel.  Output rows are the occupied voxels in
 *  ascending (ix,iy,iz) (open3d's hash-map order is unspecified); out_pts / out_normals fp64 [n,3] capacity;
 *  out_count int32 [1] (device) = number of rows, or -1 when the voxel grid has more cells than the workspace holds
 *  (the int32 at byte 56 of ws then holds the number of cells needed: call again with
 *  pdsc_workspace_bytes(PDSC_OP_VOXEL_DOWN_SAMPLE, 0, n, 0, cells, 0, 0)). */
int pdsc_voxel_down_sample(const double* pts, const double* normals, int n, double voxel, double* out_pts,
                           double* out_normals, int32_t* out_count, void* ws, size_t ws_bytes, void* stream);

/* (N1) registration.compute_fpfh_feature (demo_registration.py:41) from the hybrid-search lists of the same cloud.
 *  spfh_ws fp64 [n,33] scratch; fpfh_out fp64 [n,33] (row i = the 33 bins of point i). */
int pdsc_fpfh(const double* pts, const double* normals, int n, const int32_t* idx, const double* dist2,
              const int32_t* count, int max_nn, double* spfh_ws, double* fpfh_out, void* stream);

/* (N1) argmin_j sqrt(2 - 2 a_i.b_j + 1e-6) in float64, first minimum (demo_registration.py:101-103).
 *  A fp64 [Na,D], B fp64 [Nb,D], D <= 64; idx_out int64 [Na]; val_out (nullable) fp64 [Na] the minimum. */
int pdsc_nn_match_f64(const double* A, const double* B, int Na, int Nb, int D, int64_t* idx_out, double* val_out,
                      void* ws, size_t ws_bytes, void* stream);

/* (N4) point-to-point ICP on the key points; replaces icp_refine evaluation/benchmark_utils.py:40-56, i.e. open3d's
 *  registration_icp(source, target, max_dist, init, TransformationEstimationPointToPoint()) with
 *  ICPConvergenceCriteria(rel_fitness, rel_rmse, max_iter) (the reference uses 0.1 and the defaults 1e-6, 1e-6, 30).
 *  src, tgt fp32 [B,N,3]; trans_inout fp32 [B,4,4] (initial transform in, refined out); B independent pairs.
 *  Each round: every transformed source point is paired with its nearest target point closer than max_dist (ties to
 *  the lower index), fitness = pairs / N, rmse over the pairs; stop when both change by less than the criteria,
 *  else T <- Umeyama(pairs, no scaling) * T.  Arithmetic in float64.  Optional outputs: stats_out fp64 [B,2]
 *  (fitness, inlier rmse of the last association), iters_out int32 [B] (updates applied), corr_out int32 [B,N]
 *  (target index paired with each source point in the last association, -1 for none). */
int pdsc_icp_refine(const float* src, const float* tgt, float* trans_inout, int B, int N, double max_dist, int max_iter,
                    double rel_fitness, double rel_rmse, double* stats_out, int32_t* iters_out, int32_t* corr_out,
                    void* ws, size_t ws_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* POINTDSC_B200_H */
