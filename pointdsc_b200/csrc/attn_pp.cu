// (a) Compatibility-weighted non-local attention on the 5th-generation tensor cores, ping-pong form.
// Replaces models/PointDSC.py:150-153 + NonLocalBlock.forward :39-42.
//
//   S = Q K^T / sqrt(C)            tcgen05.mma kind::f16, fp32 accumulators in TMEM
//   W = softmax_j(compat_ij S_ij)  CUDA cores: compat recomputed from 6 floats per correspondence
//                                  (never stored), online softmax with a lazily raised reference max
//   msg = W V                      tcgen05.mma with P read from TMEM (aliasing S) and V^T from smem
//
// Precision: every fp32 operand x is split into fp16 hi + fp16 lo (22 mantissa bits), fp32 accumulation.
//   PDSC_ATTN_SPLIT16      (default) both contractions as hi*hi + hi*lo + lo*hi (3 MMAs per k-step).
//   PDSC_ATTN_SPLIT16_PV2  P V without the P_lo V_hi term, the row sum taken over the rounded P_hi.  Measured in round 2:
//                          the rounding of P (2^-12 relative, independent per key) acts like a 7e-4 perturbation of the
//                          logits; 2e-4 message error on random inputs and 3e-4 in the 12-layer confidences of the
//                          KITTI-shaped batch -- outside the tolerance -- for 6 % of the launch time.  Not the default.
//   PDSC_ATTN_FP16       hi*hi only.
//
// One CTA works on 256 query rows of one pair as two 128-row sub-tiles X = A, B that alternate on
// the tensor pipe (while the CUDA cores do the softmax of one, the tensor cores do the MMAs of the
// other), looping over 64-key tiles:
//   warps 0..15  softmax: warp w -> TMEM lane quarter w%4, sub-tile X = w/8, 16-lane half (w/4)%2.
//                Two threads per query row (lane l and l+16 of one warp, tcgen05 .16x32bx2 shape)
//                take 32 keys each, so the row maximum / sum exchange is one warp shuffle and no
//                warp ever waits for another warp inside a tile.
//   warp 16      producer: bulk-async (TMA engine) copies of K, V^T, key-point and Q_lo tiles
//   warp 17      MMA issuer (one elected lane), owns the TMEM allocation
// TMEM columns: S_A [0,64) S_B [64,128) O_A [128,256) O_B [256,384) Qhi_A [384,448) Qhi_B [448,512).
// P overwrites S in place as fp16 pairs (hi: columns 0..31, lo: 32..63 of the S block).
#include "common.cuh"
#include "sm100.cuh"

#include <math.h>
#include <algorithm>
#include <stdlib.h>

namespace pdsc {

using namespace sm100;

constexpr int PP_C = 128;                 // channels (head dim)
constexpr int PP_SUB = 128;               // rows per sub-tile
constexpr int PP_BM = 2 * PP_SUB;         // query rows per CTA
constexpr int PP_BN = 64;                 // keys per tile
constexpr uint32_t PP_QTILE_BYTES = 65536;   // per 128 rows: hi blk(ch 0-63) | hi blk(ch 64-127) | lo blk | lo blk, 16 KB swizzled blocks [128 rows x 128 B]
constexpr uint32_t PP_QLO_OFF = 32768;
constexpr uint32_t PP_QLO_BYTES = 32768;
constexpr uint32_t PP_KTILE_BYTES = 32768;   // per 64 keys: hi blk(ch 0-63) | hi blk(ch 64-127) | lo blk | lo blk, 8 KB each
constexpr uint32_t PP_KBLK_BYTES = 8192;
constexpr uint32_t PP_VTILE_BYTES = 32768;   // per 64 keys: hi [128 ch rows x 64 keys] | lo, 16 KB each
constexpr uint32_t PP_VHALF_BYTES = 16384;
// key points per key PAIR: sx0 sx1 sy0 sy1 sz0 sz1 tx0 tx1 ty0 ty1 tz0 tz1 (48 B); pairs 16..31 sit 16 B further so
// that the two half-warps (which read pairs i and 16 + i in one LDS.128) hit different banks
constexpr uint32_t PP_PTS_FLOATS = (PP_BN / 2) * 12 + 4;
constexpr uint32_t PP_PTS_BYTES = PP_PTS_FLOATS * 4;
// Lazy reference maximum (log2 units, always an integer): raised when a row maximum would put P above 2^PP_RESCALE_LOG2,
// and then set so that the maximum sits at 2^PP_TOP_LOG2.  fp16 holds P up to 2^15, so weights down to 2^-(14+6) of the
// row maximum stay normal numbers and down to 2^-30 subnormal (they were flushed 2^6 earlier with the maximum at 2^0).
constexpr float PP_RESCALE_LOG2 = 14.0f;
constexpr float PP_TOP_LOG2 = 6.0f;
constexpr uint32_t PT_S = 0, PT_O = 128, PT_Q = 384;
constexpr int PP_SOFT_WARPS = 16;
constexpr int PP_DEFAULT_CLUSTER = 1;                  // CTAs per cluster unless PDSC_ATTN_CLUSTER says otherwise
constexpr int PP_THREADS = (PP_SOFT_WARPS + 4) * 32;   // + one warpgroup of role warps (producer, MMA issuer, 2 idle)

struct PpSmem {
    uint8_t qlo[2][PP_QLO_BYTES];
    uint8_t k[2][PP_KTILE_BYTES];
    uint8_t v[2][PP_VTILE_BYTES];
    float4 pts[2][PP_PTS_FLOATS / 4 + 3];     // (+3: stages stay 64-byte aligned)
    uint64_t q_full[2], q_free;
    uint64_t k_full[2], k_empty[2], v_full[2], v_empty[2], pts_full[2], pts_empty[2];
    uint64_t s_full[2], p_full[2], pv_done[2];
    uint32_t tmem_base;
};

// byte offset of (row r, 16-byte chunk c) inside a block of 128-byte rows with the 128B swizzle
__host__ __device__ inline uint32_t pp_swz(uint32_t r, uint32_t c) { return r * 128u + ((c ^ (r & 7u)) << 4); }

__device__ __forceinline__ void pp_split8(const float (&x)[8], uint4& hi, uint4& lo) {
    uint32_t h[4], l[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        h[i] = pack_half2(x[2 * i], x[2 * i + 1]);
        float2 back = unpack_half2(h[i]);
        l[i] = pack_half2(x[2 * i] - back.x, x[2 * i + 1] - back.y);
    }
    hi = make_uint4(h[0], h[1], h[2], h[3]);
    lo = make_uint4(l[0], l[1], l[2], l[3]);
}

// ------------------------------------------------------------------------------------------------
// Prep: fp32 q,k,v [B,N,C] -> fp16 hi/lo tiles.  K, V^T and Q_lo tiles are written as the exact
// shared-memory image the MMAs read (so a tile is ONE contiguous bulk copy; V transposed so that
// keys become the contraction dim).  Q_hi is row-linear (it goes to TMEM through registers).
// q is pre-multiplied by log2(e)/sqrt(C); the key points are divided by sigma_d, so that
// compat = sat(1 - (|ds| - |dt|)^2), and packed per key pair.  One CTA per 128 rows.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) attn_pp_prep_kernel(const float* __restrict__ q, const float* __restrict__ k,
                                                           const float* __restrict__ v, int64_t ld,
                                                           const float* __restrict__ src, const float* __restrict__ tgt,
                                                           int N, int NT128, float qscale, float inv_sigma,
                                                           uint8_t* __restrict__ Qw, uint8_t* __restrict__ Kw,
                                                           uint8_t* __restrict__ Vw, float* __restrict__ Pw,
                                                           int* __restrict__ counters, int n_counters) {
    extern __shared__ __align__(16) float vt[];   // [128 keys][129] staging for the V transpose
    const int t = blockIdx.x, b = blockIdx.y;
    if (t == 0 && b == 0)                         // arrival counters of the main kernel start at zero
        for (int i = threadIdx.x; i < n_counters; i += blockDim.x) counters[i] = 0;
    const int n0 = t * PP_SUB;
    const size_t tile = (size_t)b * NT128 + t;
    uint8_t* qd = Qw + tile * PP_QTILE_BYTES;
    uint8_t* kd = Kw + tile * 2 * PP_KTILE_BYTES;   // two 64-key tiles
    uint8_t* vd = Vw + tile * 2 * PP_VTILE_BYTES;
    // Q and K: thread task = (row r, 8 consecutive channels)
    for (int task = threadIdx.x; task < 128 * 16; task += 256) {
        const int r = task >> 4, c8 = task & 15;
        const int n = n0 + r;
        float xq[8] = {0, 0, 0, 0, 0, 0, 0, 0}, xk[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (n < N) {
            const float4* pq = reinterpret_cast<const float4*>(q + ((size_t)b * N + n) * ld + c8 * 8);
            const float4* pk = reinterpret_cast<const float4*>(k + ((size_t)b * N + n) * ld + c8 * 8);
            float4 a0 = pq[0], a1 = pq[1], b0 = pk[0], b1 = pk[1];
            xq[0] = a0.x * qscale; xq[1] = a0.y * qscale; xq[2] = a0.z * qscale; xq[3] = a0.w * qscale;
            xq[4] = a1.x * qscale; xq[5] = a1.y * qscale; xq[6] = a1.z * qscale; xq[7] = a1.w * qscale;
            xk[0] = b0.x; xk[1] = b0.y; xk[2] = b0.z; xk[3] = b0.w; xk[4] = b1.x; xk[5] = b1.y; xk[6] = b1.z; xk[7] = b1.w;
        }
        uint4 hi, lo;
        pp_split8(xq, hi, lo);
        *reinterpret_cast<uint4*>(qd + (uint32_t)(c8 >> 3) * 16384u + pp_swz(r, c8 & 7)) = hi;
        *reinterpret_cast<uint4*>(qd + PP_QLO_OFF + (uint32_t)(c8 >> 3) * 16384u + pp_swz(r, c8 & 7)) = lo;
        pp_split8(xk, hi, lo);
        uint8_t* kt = kd + (size_t)(r >> 6) * PP_KTILE_BYTES + (uint32_t)(c8 >> 3) * PP_KBLK_BYTES + pp_swz(r & 63, c8 & 7);
        *reinterpret_cast<uint4*>(kt) = hi;
        *reinterpret_cast<uint4*>(kt + 2 * PP_KBLK_BYTES) = lo;
    }
    // V^T: stage the [key][channel] tile, then thread task = (channel row, 8 consecutive keys)
    for (int task = threadIdx.x; task < 128 * 32; task += 256) {
        const int r = task >> 5, c4 = task & 31;
        const int n = n0 + r;
        float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
        if (n < N) x = *reinterpret_cast<const float4*>(v + ((size_t)b * N + n) * ld + c4 * 4);
        float* d = vt + r * 129 + c4 * 4;
        d[0] = x.x; d[1] = x.y; d[2] = x.z; d[3] = x.w;
    }
    __syncthreads();
    for (int task = threadIdx.x; task < 128 * 16; task += 256) {
        const int ch = task & 127, k8 = task >> 7;   // lanes vary over channels: conflict-free smem reads
        float x[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = vt[(k8 * 8 + i) * 129 + ch];
        uint4 hi, lo;
        pp_split8(x, hi, lo);
        uint8_t* vtile = vd + (size_t)(k8 >> 3) * PP_VTILE_BYTES + pp_swz(ch, k8 & 7);
        *reinterpret_cast<uint4*>(vtile) = hi;
        *reinterpret_cast<uint4*>(vtile + PP_VHALF_BYTES) = lo;
    }
    if (threadIdx.x < PP_SUB) {
        // packed-pair SoA layout so the softmax loop can use FADD2/FFMA2 on two keys at once
        const int n = min(n0 + (int)threadIdx.x, N - 1);
        const float* s = src + ((size_t)b * N + n) * 3;
        const float* g = tgt + ((size_t)b * N + n) * 3;
        const int pr = (threadIdx.x & 63) >> 1;
        float* d = Pw + (tile * 2 + (threadIdx.x >> 6)) * PP_PTS_FLOATS + pr * 12 + (pr >= 16 ? 4 : 0) + (threadIdx.x & 1);
        d[0] = s[0] * inv_sigma; d[2] = s[1] * inv_sigma; d[4] = s[2] * inv_sigma;
        d[6] = g[0] * inv_sigma; d[8] = g[1] * inv_sigma; d[10] = g[2] * inv_sigma;
    }
}

// ------------------------------------------------------------------------------------------------
// TMEM <-> registers with the .16x32bx2 shape: the warp touches 16 TMEM lanes starting at the lane
// of `taddr`; thread t holds lane (t & 15), columns [col + (t >> 4) * SPLIT, ... + count).
// (layout verified on hardware by scripts/micro/tmem_shapes.cu)
// ------------------------------------------------------------------------------------------------
template <int SPLIT>
__device__ __forceinline__ void tmem_ld_h32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.16x32bx2.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32], %33;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr), "n"(SPLIT)
        : "memory");
}
template <int SPLIT>
__device__ __forceinline__ void tmem_st_h32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.16x32bx2.x32.b32 [%0], %33, "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
        "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
        "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]), "n"(SPLIT)
        : "memory");
}
template <int SPLIT>
__device__ __forceinline__ void tmem_st_h16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.16x32bx2.x16.b32 [%0], %17, "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "n"(SPLIT)
        : "memory");
}

// ------------------------------------------------------------------------------------------------
// MMA issue helpers (one thread).  Shared-memory operands are K-major 128B-swizzled blocks; a
// 16-element k step advances the start address by 32 bytes inside the 128-byte row.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t pp_desc_add(uint64_t desc, uint32_t byte_off) { return desc + (uint64_t)(byte_off >> 4); }
// Broadcast from lane 0: tells the compiler the value is warp-uniform (uniform registers, no per-MMA R2UR loop).
__device__ __forceinline__ uint64_t pp_uniform64(uint64_t v) {
    const uint32_t lo = __shfl_sync(0xffffffffu, (uint32_t)v, 0), hi = __shfl_sync(0xffffffffu, (uint32_t)(v >> 32), 0);
    return ((uint64_t)hi << 32) | lo;
}

// One asm statement issues a whole run of MMAs (one hi/lo term of a contraction) with a single elect.sync,
// TMEM addresses as immediates and the descriptors advanced inside the asm, so ptxas sees one straight-line
// block with warp-uniform operands.  Every lane executes the statement; the elected lane issues.
#define PP_TS "@pe tcgen05.mma.cta_group::1.kind::f16 [td], [ta], db, idc, pt;\n\t"
#define PP_SS "@pe tcgen05.mma.cta_group::1.kind::f16 [td], da, db, idc, pt;\n\t"
#define PP_ADV_TS(boff) "add.u32 ta, ta, 8;\n\tadd.s64 db, db, " #boff ";\n\t"
#define PP_ADV_SS(aoff, boff) "add.s64 da, da, " #aoff ";\n\tadd.s64 db, db, " #boff ";\n\t"
#define PP_HEAD                                                                      \
    "{\n\t.reg .pred pe, p0, pt;\n\t.reg .b64 da, db;\n\t.reg .b32 ta, td, idc;\n\t" \
    "elect.sync _|pe, 0xffffffff;\n\t"                                               \
    "setp.ne.b32 p0, %1, 0;\n\t"                                                     \
    "setp.eq.b32 pt, 0, 0;\n\t"                                                      \
    "mov.b32 td, %2;\n\tmov.b32 idc, %3;\n\tmov.b64 db, %0;\n\t"

// 8 k-steps, A from TMEM columns A0 + 8 kk, B = 64-row blocks of 8 KB (k-steps 0..3 in block 0, 4..7 in block 1)
template <uint32_t D, uint32_t A0>
__device__ __forceinline__ void pp_qk_run_ts(uint64_t k_desc, uint32_t acc_first) {
    asm volatile(PP_HEAD "mov.b32 ta, %4;\n\t"
                 "@pe tcgen05.mma.cta_group::1.kind::f16 [td], [ta], db, idc, p0;\n\t"
                 PP_ADV_TS(2) PP_TS PP_ADV_TS(2) PP_TS PP_ADV_TS(2) PP_TS
                 PP_ADV_TS(506) PP_TS PP_ADV_TS(2) PP_TS PP_ADV_TS(2) PP_TS PP_ADV_TS(2) PP_TS "}"
                 ::"l"(k_desc), "r"(acc_first), "n"(D), "n"(umma_idesc_f16(128, 64)), "n"(A0)
                 : "memory");
}
// same with A from shared memory (Q_lo: 128-row blocks of 16 KB)
template <uint32_t D>
__device__ __forceinline__ void pp_qk_run_ss(uint64_t q_desc, uint64_t k_desc, uint32_t acc_first) {
    asm volatile(PP_HEAD "mov.b64 da, %4;\n\t"
                 "@pe tcgen05.mma.cta_group::1.kind::f16 [td], da, db, idc, p0;\n\t"
                 PP_ADV_SS(2, 2) PP_SS PP_ADV_SS(2, 2) PP_SS PP_ADV_SS(2, 2) PP_SS
                 PP_ADV_SS(1018, 506) PP_SS PP_ADV_SS(2, 2) PP_SS PP_ADV_SS(2, 2) PP_SS PP_ADV_SS(2, 2) PP_SS "}"
                 ::"l"(k_desc), "r"(acc_first), "n"(D), "n"(umma_idesc_f16(128, 64)), "l"(q_desc)
                 : "memory");
}
// 4 k-steps, A = P from TMEM columns A0 + 8 kk, B = one 128-row block (V^T half)
template <uint32_t D, uint32_t A0>
__device__ __forceinline__ void pp_pv_run(uint64_t v_desc, uint32_t acc_first) {
    asm volatile(PP_HEAD "mov.b32 ta, %4;\n\t"
                 "@pe tcgen05.mma.cta_group::1.kind::f16 [td], [ta], db, idc, p0;\n\t"
                 PP_ADV_TS(2) PP_TS PP_ADV_TS(2) PP_TS PP_ADV_TS(2) PP_TS "}"
                 ::"l"(v_desc), "r"(acc_first), "n"(D), "n"(umma_idesc_f16(128, 128)), "n"(A0)
                 : "memory");
}

// S_X[128 x 64 keys] = Q K^T.  terms: bit 0 = Qhi Khi, bit 1 = Qhi Klo, bit 2 = Qlo Khi (0 = no MMAs: timing experiments).
constexpr int PP_QK_HH = 1, PP_QK_HL = 2, PP_QK_LH = 4;
// The fp32 accumulator in TMEM is updated with truncation (measured in round 1), once per MMA, at the magnitude of the
// running sum.  The two cross terms are ~2^-11 of the main term, so they go FIRST: their 16 accumulations then truncate at
// 2^-11 of the logit's magnitude and only the 8 hi*hi accumulations at the full one (24 before: about a third of the bias).
template <int X>
__device__ __forceinline__ void pp_issue_qk(uint64_t qlo_desc, uint64_t k_desc, int terms) {
    if (!(terms & PP_QK_HH)) return;
    uint32_t acc = 0u;
    if (terms & PP_QK_LH) { pp_qk_run_ss<PT_S + 64u * X>(qlo_desc, k_desc, acc); acc = 1u; }
    if (terms & PP_QK_HL) { pp_qk_run_ts<PT_S + 64u * X, PT_Q + 64u * X>(pp_desc_add(k_desc, 2u * PP_KBLK_BYTES), acc); acc = 1u; }
    pp_qk_run_ts<PT_S + 64u * X, PT_Q + 64u * X>(k_desc, acc);
}
// O_X[128 x 128 ch] (+)= P[128 x 64 keys] V.  terms = 3: Phi Vhi + Phi Vlo + Plo Vhi; 2: Phi Vhi + Phi Vlo; 1: Phi Vhi.
template <int X>
__device__ __forceinline__ void pp_issue_pv(uint64_t v_desc, int terms, uint32_t acc) {
    if (terms <= 0) return;
    pp_pv_run<PT_O + 128u * X, PT_S + 64u * X>(v_desc, acc);
    if (terms > 1) pp_pv_run<PT_O + 128u * X, PT_S + 64u * X>(pp_desc_add(v_desc, PP_VHALF_BYTES), 1u);
    if (terms > 2) pp_pv_run<PT_O + 128u * X, PT_S + 64u * X + 32u>(v_desc, 1u);
}

__device__ __forceinline__ float pp_fma_sat(float a, float b, float c) {
    float d;
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

// (Round 2 measured the alternative c = (a - b) * rsqrt(a + b + 2 sqrt(a b)) with a = |ds|^2, b = |dt|^2 -- same two MUFU
// operations, no error proportional to the distance: +7 % launch time, no change of the confidence error on the
// KITTI-shaped batch (profiles/r02_kitti_error_budget.txt), so the direct form below stays.)
// compat of this thread's query point against NP key pairs -> w[NP] (two keys per float2)
template <int NP>
__device__ __forceinline__ void pp_compat_pairs(const float4* pts, const float2 (&qp)[6], float2 (&w)[NP]) {
#pragma unroll
    for (int i = 0; i < NP; ++i) {
        const float4 pa = pts[3 * i], pb = pts[3 * i + 1], pc = pts[3 * i + 2];
        float2 dx = __fadd2_rn(qp[0], make_float2(-pa.x, -pa.y));
        float2 dy = __fadd2_rn(qp[1], make_float2(-pa.z, -pa.w));
        float2 dz = __fadd2_rn(qp[2], make_float2(-pb.x, -pb.y));
        float2 a2 = __ffma2_rn(dz, dz, __ffma2_rn(dy, dy, __fmul2_rn(dx, dx)));
        const float2 ds = make_float2(fast_sqrt(a2.x), fast_sqrt(a2.y));
        dx = __fadd2_rn(qp[3], make_float2(-pb.z, -pb.w));
        dy = __fadd2_rn(qp[4], make_float2(-pc.x, -pc.y));
        dz = __fadd2_rn(qp[5], make_float2(-pc.z, -pc.w));
        a2 = __ffma2_rn(dz, dz, __ffma2_rn(dy, dy, __fmul2_rn(dx, dx)));
        const float2 dt = make_float2(fast_sqrt(a2.x), fast_sqrt(a2.y));
        const float2 c = __fadd2_rn(ds, make_float2(-dt.x, -dt.y));
        w[i] = make_float2(pp_fma_sat(-c.x, c.x, 1.0f), pp_fma_sat(-c.y, c.y, 1.0f));   // max(0, 1 - c^2)
    }
}

// ------------------------------------------------------------------------------------------------
// Work decomposition ("stream-K"): the (pair, 256-row block, 64-key tile) steps are linearised and
// cut into `gridDim.x` equal contiguous ranges, so every SM gets the same number of steps whatever
// B and N are.  A range is walked as segments = (block, key tiles [j0,j1)).  A segment covering all
// key tiles of its block writes the message directly; otherwise the CTA stores its un-normalised
// partial (O, m, l) and the LAST contributor of that block (atomic counter) combines them.
// A CTA has at most two partial segments (its first and its last) -> two slots per CTA.
// Chain limit: tcgen05 accumulates O in fp32 with truncation, so the error of one accumulation chain grows linearly with
// the number of key tiles folded into the TMEM accumulator (measured on the N=10000 KITTI-shaped case: confidences drift
// by 1.5e-4 when whole 157-tile blocks are accumulated in one go, 6e-5 with <= 48-tile chains).  For long blocks the
// softmax threads therefore store O (and its reference max) every `chain` tiles of a segment into a slot of the workspace
// (L2-resident, stores only) and the next PV MMA starts a fresh chain (accumulate = 0); the segment epilogue sums the
// slots and the last chain with properly rounded fp32 adds.  No pipeline drain, no extra partials.  Production launch
// shape only (the cluster-of-2 and debug instantiations run without the limit).
// ------------------------------------------------------------------------------------------------
struct PpParams {
    const uint8_t* Qw; const uint8_t* Kw; const uint8_t* Vw; const float* Pw;
    float* out; int64_t ld_out;
    float* part_o;          // [grid][2 slots][2 sub-tiles][64 KB in the owner threads' order, see pp_priv_index]
    float* part_ml;         // [grid][2][2][128][2]  (m, l)
    float* chain_o;         // [grid][PP_CHAIN_SLOTS][2 sub-tiles][64 KB, pp_priv_index]: O of the finished chains of the current segment
    float* chain_m;         // [grid][PP_CHAIN_SLOTS][2][128]: the reference max each was stored with
    int chain;              // key tiles per accumulation chain (0 = no limit)
    int* counters;          // [B * NRBp][2]
    int N, NT;              // rows, 64-key tiles per pair
    int NRBp;               // 256-row blocks per pair rounded up to even: tile / counter indexing
    int NU;                 // work units per pair: 256-row blocks (no cluster) or 512-row block pairs (cluster of 2)
    long long total_steps;  // B * NU * NT
    int qk_terms, pv_terms, dbg;   // hi/lo product terms issued per contraction
    uint32_t hint_soft, hint_role; // suspend-time hints (ns) of the barrier waits of the softmax warps / the role warps
    long long* trace;       // optional timeline of CTA 0 (PDSC_ATTN_TRACE): trace[step * 16 + event] = SM clock
};

constexpr int PP_CHAIN_SLOTS = 8;     // finished chains per segment that the workspace holds
struct PpSeg { int blk, j0, j1; };
// rows 0..239: per key-tile step; rows 240..255: per segment (0/1 segment start A/B, 2/3 body of the segment done, 4/5 last PV seen, 6/7 epilogue done)
#define PP_TRACE(step, ev) do { if (tracing && (step) < 256) p.trace[(step) * 16 + (ev)] = clock64(); } while (0)

__device__ __forceinline__ long long pp_range_lo(long long total, int c, int G) { return total * c / G; }

// si-th segment of the range [lo, hi); returns false past the end.  `pos` carries the walk state.
__device__ __forceinline__ bool pp_next_segment(long long& pos, long long hi, int NT, PpSeg& s) {
    if (pos >= hi) return false;
    s.blk = (int)(pos / NT);
    s.j0 = (int)(pos - (long long)s.blk * NT);
    const long long left = hi - pos;
    s.j1 = (left < (long long)(NT - s.j0)) ? s.j0 + (int)left : NT;
    pos += s.j1 - s.j0;
    return true;
}

__device__ __forceinline__ float pp_ldcg(const float* p) {
    float v;
    asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ float4 pp_ldcg4(const float* p) {
    float4 v;
    asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

// Private order of a sub-tile's O (128 rows x 128 channels) in the workspace (stream-K partials, finished chains): these
// buffers are written and read back by the thread with the same (warp of the sub-tile, lane), which owns 64 channels of one
// row as 16 float4 (f = 8 * chunk + i).  Stored as [warp][f][lane] every warp instruction moves 512 contiguous bytes; stored
// row-major (each lane 128 bytes from its neighbour, as until round 2) it touched 32 lines and the epilogue of a segment
// spent 14 k clocks in the LSU (profiles/experiments/r02_attn_segment_boundaries.txt).
__device__ __forceinline__ uint32_t pp_priv_index(int w8, int f, int lane) { return (uint32_t)((w8 * 16 + f) * 32 + lane); }

// In-register transpose of an 8 x 8 matrix of float4 over the 8 lanes of a group (k = lane & 7): before, lane k holds the
// eight float4 of ITS 128-byte piece; after, it holds float4 number k of the pieces of the lanes 8g .. 8g+7.  A warp store of
// element j then writes four whole 128-byte lines (one per group) instead of 16 bytes into each of 32 lines.
__device__ __forceinline__ void pp_transpose8(uint32_t (&r)[32], int k) { warp_transpose8(r, k); }      // (sm100.cuh)

// Rows of the output: chunk c (32 channels) of the 64 channels every lane of the warp owns, already scaled.  `tile_row0` =
// first row of the warp's 16 rows in the pair, lanes 0-15 own channels [0,64) of those rows and lanes 16-31 channels [64,128).
__device__ __forceinline__ void pp_store_rows(float* out_pair, int64_t ld_out, int tile_row0, int N, int c, uint32_t (&r)[32], int lane) {
    const int g = lane >> 3, k = lane & 7;
    pp_transpose8(r, k);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int src_lane = 8 * g + j;
        const int grow = tile_row0 + (src_lane & 15);
        if (grow < N)
            *reinterpret_cast<uint4*>(out_pair + (size_t)grow * ld_out + (src_lane >> 4) * 64 + c * 32 + k * 4) =
                make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
    }
}

// ------------------------------------------------------------------------------------------------
// Main kernel.
// ------------------------------------------------------------------------------------------------
// CL = CTAs per cluster.  CL == 2: the two CTAs of a cluster walk the same key tiles of one pair for two neighbouring
// 256-row blocks; each loads HALF of every K / V^T tile and multicasts it to both (cp.async.bulk .multicast::cluster),
// which halves the L2 -> SM traffic; a stage is released by both MMA issuers (multicast tcgen05.commit).
// DBG: instantiation with the timing-experiment knobs (PDSC_ATTN_DBG) and the timeline trace compiled in.
// PV3: the P_lo V_hi term is issued (PDSC_ATTN_SPLIT16), so the softmax threads also store the fp16 remainder of P.
template <int CL, bool DBG, bool CHAIN = false, bool PV3 = false>
__global__ void __launch_bounds__(PP_THREADS, 1) attn_pp_kernel(const PpParams p) {
    constexpr int NP = 16;                   // key pairs per thread per tile (32 keys)
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // the 128B swizzle is anchored at 1024-byte boundaries of the shared window
    PpSmem& sm = *reinterpret_cast<PpSmem*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
    __shared__ int last_flag[2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int G = gridDim.x / CL, cta = blockIdx.x;        // G work ranges; range index = cluster index
    const int rank = CL == 1 ? 0 : (int)cluster_ctarank();
    const int rng = blockIdx.x / CL;
    const int NT = p.NT;
    const long long lo = pp_range_lo(p.total_steps, rng, G), hi = pp_range_lo(p.total_steps, rng + 1, G);
    const bool tracing = DBG && p.trace != nullptr && cta == 0 && lane == 0;
    constexpr uint16_t CL_MASK = (uint16_t)((1u << CL) - 1u);

    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(&sm.q_full[i], PP_SOFT_WARPS / 2 + 1);
            mbar_init(&sm.k_full[i], 1); mbar_init(&sm.k_empty[i], CL);     // released by the MMA issuer of every CTA in the cluster
            mbar_init(&sm.v_full[i], 1); mbar_init(&sm.v_empty[i], CL);
            mbar_init(&sm.pts_full[i], 1); mbar_init(&sm.pts_empty[i], PP_SOFT_WARPS);
            mbar_init(&sm.s_full[i], 1); mbar_init(&sm.p_full[i], PP_SOFT_WARPS / 2); mbar_init(&sm.pv_done[i], 1);
        }
        mbar_init(&sm.q_free, 1);
        fence_barrier_init();
    }
    if (warp == PP_SOFT_WARPS + 1) tmem_alloc(&sm.tmem_base, 512);
    tc_fence_before();
    __syncthreads();
    if (CL > 1) cluster_sync_all();            // every CTA's barriers exist before a peer signals them
    tc_fence_after();
    // The allocation is the whole 512-column TMEM of the SM (one CTA per SM), so its base is column 0, lane 0:
    // every TMEM address below is a compile-time constant.
    if (sm.tmem_base != 0u) __trap();
    constexpr uint32_t tmem = 0u;

    if (warp >= PP_SOFT_WARPS) {
        // role warpgroup: give registers back to the softmax warpgroups
        asm volatile("setmaxnreg.dec.sync.aligned.u32 32;");
    }
    if (warp == PP_SOFT_WARPS) {
        // ============================== producer ==============================
        if (lane == 0) {
            long long pos = lo;
            PpSeg s;
            uint32_t kt = 0;
            for (int si = 0; pp_next_segment(pos, hi, NT, s); ++si) {
                const int b = s.blk / p.NU, rb = (s.blk - b * p.NU) * CL + rank;
                const size_t q128 = (size_t)b * (2 * p.NRBp) + 2 * rb;     // first 128-row tile of the block
                const size_t kt0 = (size_t)b * (4 * p.NRBp);               // first 64-key tile of the pair
                if (si > 0) mbar_wait_role(&sm.q_free, (si - 1) & 1);      // all MMAs of the previous segment are done
                for (int x = 0; x < 2; ++x) {
                    mbar_arrive_expect_tx(&sm.q_full[x], PP_QLO_BYTES);
                    bulk_g2s(sm.qlo[x], p.Qw + (q128 + x) * PP_QTILE_BYTES + PP_QLO_OFF, PP_QLO_BYTES, &sm.q_full[x]);
                }
                for (int j = s.j0; j < s.j1; ++j, ++kt) {
                    const int st = kt & 1;
                    const uint32_t par = ((kt >> 1) - 1) & 1;               // previous use of this stage
                    if (kt >= 2) mbar_wait_role(&sm.k_empty[st], par);
                    mbar_arrive_expect_tx(&sm.k_full[st], PP_KTILE_BYTES);
                    if (CL == 1) {
                        bulk_g2s(sm.k[st], p.Kw + (kt0 + j) * PP_KTILE_BYTES, PP_KTILE_BYTES, &sm.k_full[st]);
                    } else {
                        constexpr uint32_t part = PP_KTILE_BYTES / CL;
                        bulk_g2s_multicast(sm.k[st] + rank * part, p.Kw + (kt0 + j) * PP_KTILE_BYTES + rank * part, part, &sm.k_full[st], CL_MASK);
                    }
                    PP_TRACE(kt, 0);
                    if (kt >= 2) mbar_wait_role(&sm.pts_empty[st], par);
                    mbar_arrive_expect_tx(&sm.pts_full[st], PP_PTS_BYTES);
                    bulk_g2s(sm.pts[st], p.Pw + (kt0 + j) * PP_PTS_FLOATS, PP_PTS_BYTES, &sm.pts_full[st]);
                    if (kt >= 2) mbar_wait_role(&sm.v_empty[st], par);
                    mbar_arrive_expect_tx(&sm.v_full[st], PP_VTILE_BYTES);
                    if (CL == 1) {
                        bulk_g2s(sm.v[st], p.Vw + (kt0 + j) * PP_VTILE_BYTES, PP_VTILE_BYTES, &sm.v_full[st]);
                    } else {
                        constexpr uint32_t part = PP_VTILE_BYTES / CL;
                        bulk_g2s_multicast(sm.v[st] + rank * part, p.Vw + (kt0 + j) * PP_VTILE_BYTES + rank * part, part, &sm.v_full[st], CL_MASK);
                    }
                    PP_TRACE(kt, 1);
                }
            }
        }
    } else if (warp == PP_SOFT_WARPS + 1) {
        // ============================== MMA issuer ==============================
        // The whole warp runs this loop (warp-uniform control flow keeps the descriptors in uniform
        // registers); only the tcgen05 instructions themselves are predicated on an elected lane.
        const uint64_t k_desc[2] = {umma_desc_sw128_kmajor(smem_u32(sm.k[0])), umma_desc_sw128_kmajor(smem_u32(sm.k[1]))};
        const uint64_t v_desc[2] = {umma_desc_sw128_kmajor(smem_u32(sm.v[0])), umma_desc_sw128_kmajor(smem_u32(sm.v[1]))};
        const uint64_t q_desc0 = pp_uniform64(umma_desc_sw128_kmajor(smem_u32(sm.qlo[0])));
        const uint64_t q_desc1 = pp_uniform64(umma_desc_sw128_kmajor(smem_u32(sm.qlo[1])));
        const int qk_terms = p.qk_terms, pv_terms = p.pv_terms;
        const uint32_t hr = p.hint_role;
        long long pos = lo;
        PpSeg s;
        uint32_t kt = 0;
        for (int si = 0; pp_next_segment(pos, hi, NT, s); ++si) {
            const int n = s.j1 - s.j0;
            // S_A(0), S_B(0)
            mbar_wait_uniform(&sm.k_full[kt & 1], (kt >> 1) & 1, hr);
            {
                const uint64_t kd = pp_uniform64(k_desc[kt & 1]);
                mbar_wait_uniform(&sm.q_full[0], si & 1, hr);
                tc_fence_after();
                pp_issue_qk<0>(q_desc0, kd, qk_terms);
                umma_commit_elect(&sm.s_full[0]);
                mbar_wait_uniform(&sm.q_full[1], si & 1, hr);
                tc_fence_after();
                pp_issue_qk<1>(q_desc1, kd, qk_terms);
                umma_commit_elect(&sm.s_full[1]);
                if (CL == 1) umma_commit_elect(&sm.k_empty[kt & 1]); else umma_commit_elect_multicast(&sm.k_empty[kt & 1], CL_MASK);
            }
            int next_chain = CHAIN ? p.chain : 0;     // first tile of the next accumulation chain
            for (int t = 0; t < n; ++t, ++kt) {
                const int st = kt & 1, nx = st ^ 1;
                const bool more = t + 1 < n;
                const uint64_t vd = pp_uniform64(v_desc[st]), kd = pp_uniform64(k_desc[nx]);
                if (CHAIN && t == next_chain) next_chain += p.chain;
                const uint32_t acc = (t == 0 || (CHAIN && t + p.chain == next_chain)) ? 0u : 1u;     // a fresh chain overwrites O
                // ---- sub-tile A ----
                mbar_wait_uniform(&sm.p_full[0], kt & 1, hr);
                PP_TRACE(kt, 2);
                mbar_wait_uniform(&sm.v_full[st], (kt >> 1) & 1, hr);
                PP_TRACE(kt, 12);
                tc_fence_after();
                pp_issue_pv<0>(vd, pv_terms, acc);
                umma_commit_elect(&sm.pv_done[0]);
                if (more) {
                    mbar_wait_uniform(&sm.k_full[nx], ((kt + 1) >> 1) & 1, hr);
                    tc_fence_after();
                    PP_TRACE(kt, 13);
                    pp_issue_qk<0>(q_desc0, kd, qk_terms);
                    umma_commit_elect(&sm.s_full[0]);
                }
                PP_TRACE(kt, 4);
                // ---- sub-tile B ----
                mbar_wait_uniform(&sm.p_full[1], kt & 1, hr);
                PP_TRACE(kt, 3);
                tc_fence_after();
                pp_issue_pv<1>(vd, pv_terms, acc);
                umma_commit_elect(&sm.pv_done[1]);
                if (CL == 1) umma_commit_elect(&sm.v_empty[st]); else umma_commit_elect_multicast(&sm.v_empty[st], CL_MASK);
                if (more) {
                    pp_issue_qk<1>(q_desc1, kd, qk_terms);
                    umma_commit_elect(&sm.s_full[1]);
                    if (CL == 1) umma_commit_elect(&sm.k_empty[nx]); else umma_commit_elect_multicast(&sm.k_empty[nx], CL_MASK);
                }
                PP_TRACE(kt, 5);
            }
            umma_commit_elect(&sm.q_free);
        }
    } else if (warp < PP_SOFT_WARPS) {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 112;");
        // ============================== softmax warps ==============================
        const int quarter = warp & 3;                 // TMEM lane quarter this warp may access
        const int x = warp >> 3;                      // sub-tile A / B
        const int lh = (warp >> 2) & 1;               // which 16 lanes of the quarter
        const int half = lane >> 4;                   // which half of the columns this thread takes
        const int w8 = warp & 7;                      // warp inside the sub-tile (order of the private workspace buffers)
        const int row = quarter * 32 + lh * 16 + (lane & 15);          // query row inside the sub-tile
        const uint32_t lane_addr = (uint32_t)(quarter * 32 + lh * 16) << 16;
        const uint32_t s_x = tmem + lane_addr + PT_S + 64u * x;
        const uint32_t o_x = tmem + lane_addr + PT_O + 128u * x;
        const uint32_t q_x = tmem + lane_addr + PT_Q + 64u * x;
        const int kp0 = half * NP;                    // first key pair of this thread inside a tile
        const int dbg = DBG ? p.dbg : 0;
        constexpr bool pv3 = PV3;
        const int N = p.N;
        const uint32_t hs = p.hint_soft;
        long long pos = lo;
        PpSeg s, s_next;
        uint32_t kt = 0;
        // Q_hi rows of a segment's block -> TMEM (this thread: 32 of the 64 words of its row), then q_full.  (Handing the next
        // block's Q_hi over BEFORE the epilogue of its predecessor, with q_free raised after the last Q K^T, was measured: 1-5 %
        // slower, profiles/experiments/r02_attn_early_q_handover.diff; reading whole rows per 8-lane group + the transpose of
        // pp_transpose8 instead of a row per lane: no gain, same file.)
        auto load_q_hi = [&](const PpSeg& sg) {
            const int b = sg.blk / p.NU, rb = (sg.blk - b * p.NU) * CL + rank;
            const size_t q128 = (size_t)b * (2 * p.NRBp) + 2 * rb + x;
            const uint8_t* qsrc = p.Qw + q128 * PP_QTILE_BYTES + (size_t)half * 16384u;   // block `half` = channels 64 half ..
            uint32_t r[32];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint4 v4 = __ldg(reinterpret_cast<const uint4*>(qsrc + pp_swz((uint32_t)row, (uint32_t)i)));
                r[4 * i] = v4.x; r[4 * i + 1] = v4.y; r[4 * i + 2] = v4.z; r[4 * i + 3] = v4.w;
            }
            tmem_st_h32<32>(q_x, r);
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.q_full[x]);
        };
        bool have = pp_next_segment(pos, hi, NT, s);
        for (int si = 0; have; ++si) {
            load_q_hi(s);
            const int b = s.blk / p.NU, rb = (s.blk - b * p.NU) * CL + rank;
            const size_t kt0 = (size_t)b * (4 * p.NRBp);
            const int grow = (2 * rb + x) * PP_SUB + row;
            const int qrow = min(grow, N - 1);
            const int warp_row0 = (2 * rb + x) * PP_SUB + quarter * 32 + lh * 16;      // first of this warp's 16 rows in the pair
            float* out_pair = p.out + (size_t)b * N * p.ld_out;
            if ((warp & 7) == 0) PP_TRACE(240 + min(si, 15), 0 + x);
            if (si == 1) PP_TRACE(224 + warp, 1);         // rows 224..239: every softmax warp at the first boundary of the CTA
            float2 qp[6];
            {
                const int qpr = (qrow & 63) >> 1;
                const float* pp = p.Pw + (kt0 + (qrow >> 6)) * PP_PTS_FLOATS + qpr * 12 + (qpr >= 16 ? 4 : 0) + (qrow & 1);
#pragma unroll
                for (int i = 0; i < 6; ++i) qp[i] = make_float2(pp[2 * i], pp[2 * i]);
            }
            float m = 0.f, l = 0.f;                   // reference max (log2 units; compat = 0 keys sit at 0) and row sum
            const int n = s.j1 - s.j0;
            int t = 0, t_chain = 0;                   // t_chain: first tile of the current accumulation chain
            do {
            const int t_end = CHAIN ? min(n, t_chain + p.chain) : n;
            for (; t < t_end; ++t, ++kt) {
                const int st = kt & 1;
                const int j = s.j0 + t;
                // compatibilities of this thread's row against its 32 keys (needs only the points)
                float2 e[NP];
                mbar_wait_role(&sm.pts_full[st], (kt >> 1) & 1, hs);
                // (the never-taken constant branch also fences the compat block: without it ptxas spreads the
                // compat arithmetic over the following waits and the kernel ran ~6% slower in round 1; re-measured in round 2
                // against three cheaper fences -- zeros, a trap after the block, none: 0.653 vs 0.660 / 0.661 / 0.665 ms)
                if (DBG ? (dbg & 1) : (N < 0)) {
#pragma unroll
                    for (int i = 0; i < NP; ++i) e[i] = make_float2(0.5f, 0.25f);
                } else {
                    pp_compat_pairs<NP>(&sm.pts[st][kp0 * 3 + half], qp, e);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&sm.pts_empty[st]);
                if ((warp & 7) == 0) PP_TRACE(kt, 6 + x);
                if (si == 1 && t == 0) PP_TRACE(224 + warp, 2);
                mbar_wait_role(&sm.s_full[x], kt & 1, hs);
                if ((warp & 7) == 0) PP_TRACE(kt, 8 + x);
                if (si == 1 && t == 0) PP_TRACE(224 + warp, 3);
                tc_fence_after();
                // e = compat * S - m  (log2 units), and its maximum over this row's keys
                float emax = -INFINITY;
                {
                    uint32_t r[32];
                    tmem_ld_h32<32>(s_x, r);
                    tmem_ld_wait();
                    const float2 neg_m = make_float2(-m, -m);
#pragma unroll
                    for (int i = 0; i < NP; ++i) {
                        const float2 ev = __ffma2_rn(e[i], make_float2(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1])), neg_m);
                        e[i] = ev;
                        emax = fmaxf(emax, fmaxf(ev.x, ev.y));
                    }
                }
                const int nvalid = N - j * PP_BN;      // keys of this tile that exist
                if (nvalid < PP_BN) {                  // last, partial tile: keys past N do not exist
                    emax = -INFINITY;
#pragma unroll
                    for (int i = 0; i < NP; ++i) {
                        const int key = half * 32 + 2 * i;
                        if (key >= nvalid) e[i].x = -INFINITY;
                        if (key + 1 >= nvalid) e[i].y = -INFINITY;
                        emax = fmaxf(emax, fmaxf(e[i].x, e[i].y));
                    }
                }
                emax = fmaxf(emax, __shfl_xor_sync(0xffffffffu, emax, 16));   // the other half of the row
                // Lazy reference max (an integer, see PP_RESCALE_LOG2): raised only when the row maximum would exceed
                // 2^14; on the first tile it is also lowered for rows whose logits are all below zero.
                const bool need = (emax > PP_RESCALE_LOG2) || (t == 0 && emax < -2.0f && emax > -INFINITY);
                if (__any_sync(0xffffffffu, need)) {
                    const float shift = need ? rintf(emax) - PP_TOP_LOG2 : 0.f;
                    const float f = fast_exp2(-shift);
                    m += shift;
                    l *= f;
                    const float2 ns = make_float2(-shift, -shift);
#pragma unroll
                    for (int i = 0; i < NP; ++i) e[i] = __fadd2_rn(e[i], ns);
                    if (t > t_chain) {                 // (the first PV of a chain overwrites O: nothing to rescale yet)
                        mbar_wait_role(&sm.pv_done[x], (kt - 1) & 1, hs);      // PV(t-1) has finished accumulating into O
                        tc_fence_after();
#pragma unroll
                        for (int c = 0; c < 2; ++c) {
                            uint32_t r[32];
                            tmem_ld_h32<64>(o_x + 32u * c, r);
                            tmem_ld_wait();
#pragma unroll
                            for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * f);
                            tmem_st_h32<64>(o_x + 32u * c, r);
                        }
                    }
                }
                // P = exp2(e) as fp16 pairs, two keys per TMEM column: hi at S columns [0,32), and (when the
                // P_lo V_hi term is issued) the fp16 remainder lo at [32,64).  Without P_lo the row sum is
                // taken over the ROUNDED weights, so numerator and denominator use the same weights.
                float2 psum = make_float2(0.f, 0.f);
                if (pv3) {
                    uint32_t phi[NP], plo[NP];
#pragma unroll
                    for (int i = 0; i < NP; ++i) {
                        const float2 ev = e[i];
                        const float2 pr = (dbg & 2) ? make_float2(1.0f + 1e-3f * ev.x, 1.0f + 1e-3f * ev.y)
                                                    : make_float2(fast_exp2(ev.x), fast_exp2(ev.y));
                        psum = __fadd2_rn(psum, pr);
                        phi[i] = pack_half2(pr.x, pr.y);
                        const float2 back = unpack_half2(phi[i]);
                        const float2 rem = __fadd2_rn(pr, make_float2(-back.x, -back.y));
                        plo[i] = pack_half2(rem.x, rem.y);
                    }
                    tmem_st_h16<16>(s_x, phi);
                    tmem_st_h16<16>(s_x + 32u, plo);
                } else {
                    uint32_t phi[NP];
#pragma unroll
                    for (int i = 0; i < NP; ++i) {
                        const float2 ev = e[i];
                        const float2 pr = (dbg & 2) ? make_float2(1.0f + 1e-3f * ev.x, 1.0f + 1e-3f * ev.y)
                                                    : make_float2(fast_exp2(ev.x), fast_exp2(ev.y));
                        phi[i] = pack_half2(pr.x, pr.y);
                        psum = __fadd2_rn(psum, unpack_half2(phi[i]));
                    }
                    tmem_st_h16<16>(s_x, phi);
                }
                l += psum.x + psum.y;
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&sm.p_full[x]);
                if ((warp & 7) == 0) PP_TRACE(kt, 10 + x);
                if (si == 1 && t == 0) PP_TRACE(224 + warp, 4);
                if (si == 0 && t == n - 1) PP_TRACE(224 + warp, 7);
            }
                if (CHAIN && t < n) {
                    // Chain limit: the finished chain's O and its reference max go to this thread's slot of the workspace (stores
                    // only: nothing waits for them); the first PV of the next chain overwrites O (accumulate = 0).
                    mbar_wait_role(&sm.pv_done[x], (kt - 1) & 1, hs);          // PV(t-1) has finished accumulating into O
                    tc_fence_after();
                    const size_t slot = ((size_t)cta * PP_CHAIN_SLOTS + (t / p.chain - 1)) * 2 + x;
                    uint4* ca = reinterpret_cast<uint4*>(p.chain_o + slot * (PP_SUB * PP_C));
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        uint32_t r[32];
                        tmem_ld_h32<64>(o_x + 32u * c, r);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            __stcg(ca + pp_priv_index(w8, c * 8 + i, lane), make_uint4(r[4 * i], r[4 * i + 1], r[4 * i + 2], r[4 * i + 3]));
                    }
                    if (half == 0) __stcg(p.chain_m + slot * PP_SUB + row, m);
                    t_chain = t;
                }
            } while (CHAIN && t < n);
            // ---- epilogue of the segment ----
            const bool have_next = pp_next_segment(pos, hi, NT, s_next);
            if ((warp & 7) == 0) PP_TRACE(240 + min(si, 15), 2 + x);
            l += __shfl_xor_sync(0xffffffffu, l, 16);
            mbar_wait_role(&sm.pv_done[x], (kt - 1) & 1, hs);
            tc_fence_after();
            if ((warp & 7) == 0) PP_TRACE(240 + min(si, 15), 4 + x);
            if (si == 0) PP_TRACE(224 + warp, 5);
            // O of the segment = the last chain (TMEM) + the finished chains (workspace slots, each relative to the reference
            // max it was stored with)
            const int nflush = CHAIN ? (n - 1) / p.chain : 0;
            auto add_chains = [&](uint32_t (&r)[32], int c) {
                if (CHAIN) {
                    for (int k = 0; k < nflush; ++k) {
                        const size_t slot = ((size_t)cta * PP_CHAIN_SLOTS + k) * 2 + x;
                        const float4* ca = reinterpret_cast<const float4*>(p.chain_o + slot * (PP_SUB * PP_C));
                        float4 a[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) a[i] = __ldcg(ca + pp_priv_index(w8, c * 8 + i, lane));     // eight loads in flight
                        const float f = exp2f(__ldcg(p.chain_m + slot * PP_SUB + row) - m);
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            r[4 * i] = __float_as_uint(__uint_as_float(r[4 * i]) + f * a[i].x);
                            r[4 * i + 1] = __float_as_uint(__uint_as_float(r[4 * i + 1]) + f * a[i].y);
                            r[4 * i + 2] = __float_as_uint(__uint_as_float(r[4 * i + 2]) + f * a[i].z);
                            r[4 * i + 3] = __float_as_uint(__uint_as_float(r[4 * i + 3]) + f * a[i].w);
                        }
                    }
                }
            };
            if (s.j0 == 0 && s.j1 == NT) {
                // all keys seen: msg = O / l, this thread writes 64 channels of its row
                const float inv_l = 1.0f / l;
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    uint32_t r[32];
                    tmem_ld_h32<64>(o_x + 32u * c, r);
                    tmem_ld_wait();
                    add_chains(r, c);
#pragma unroll
                    for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * inv_l);
                    pp_store_rows(out_pair, p.ld_out, warp_row0, N, c, r, lane);
                }
            } else {
                // partial: store (O, m, l); the last contributor of the block combines all partials
                const size_t pslot = (size_t)(cta * 2 + (si == 0 ? 0 : 1)) * 2 + x;      // cta = rng * CL + rank
                const size_t base = pslot * PP_SUB + row;
                uint4* po = reinterpret_cast<uint4*>(p.part_o + pslot * (PP_SUB * PP_C));
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    uint32_t r[32];
                    tmem_ld_h32<64>(o_x + 32u * c, r);
                    tmem_ld_wait();
                    add_chains(r, c);
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        po[pp_priv_index(w8, c * 8 + i, lane)] = make_uint4(r[4 * i], r[4 * i + 1], r[4 * i + 2], r[4 * i + 3]);
                }
                if (half == 0) {
                    p.part_ml[base * 2] = m;
                    p.part_ml[base * 2 + 1] = l;
                }
                __threadfence();
                named_bar_sync(1 + x, 256);
                if ((warp & 7) == 0 && lane == 0) last_flag[x] = atomicAdd(&p.counters[((size_t)b * p.NRBp + rb) * 2 + x], 1);
                named_bar_sync(1 + x, 256);
                const long long T = p.total_steps, x0 = (long long)s.blk * NT;
                const int c_first = (int)(((x0 + 1) * G + T - 1) / T) - 1;
                const int c_last = (int)(((x0 + NT) * G + T - 1) / T) - 1;
                if (last_flag[x] == c_last - c_first) {
                    // last arrival: leave the counter at zero for the next launch on this workspace
                    if ((warp & 7) == 0 && lane == 0) p.counters[((size_t)b * p.NRBp + rb) * 2 + x] = 0;
                    __threadfence();
                    // contributor c wrote slot 0 if the block is the first segment of its range, else slot 1: only c_first can
                    // have started before the block
                    const int first_slot = pp_range_lo(T, c_first, G) >= x0 ? 0 : 1;
                    auto part_slot = [&](int c) { return (size_t)((c * CL + rank) * 2 + (c == c_first ? first_slot : 0)) * 2 + x; };
                    if (c_last == c_first + 1) {
                        // two contributors (the usual case): every load of a pass is in flight at once; same arithmetic,
                        // term by term, as the general loop below
                        const size_t ps0 = part_slot(c_first), ps1 = part_slot(c_last);
                        const float* ml0 = p.part_ml + (ps0 * PP_SUB + row) * 2;
                        const float* ml1 = p.part_ml + (ps1 * PP_SUB + row) * 2;
                        const float m0 = pp_ldcg(ml0), l0 = pp_ldcg(ml0 + 1), m1 = pp_ldcg(ml1), l1 = pp_ldcg(ml1 + 1);
                        const float M = fmaxf(fmaxf(-INFINITY, m0), m1);
                        const float f0 = exp2f(m0 - M), f1 = exp2f(m1 - M);
                        float L = 0.f;
                        L += f0 * l0;
                        L += f1 * l1;
                        const float inv_l = 1.0f / L;
                        const float* o0 = p.part_o + ps0 * (PP_SUB * PP_C);
                        const float* o1 = p.part_o + ps1 * (PP_SUB * PP_C);
#pragma unroll 1
                        for (int ch = 0; ch < 2; ++ch) {
                            float4 a0[8], a1[8];
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                a0[i] = pp_ldcg4(o0 + 4 * (size_t)pp_priv_index(w8, ch * 8 + i, lane));
                                a1[i] = pp_ldcg4(o1 + 4 * (size_t)pp_priv_index(w8, ch * 8 + i, lane));
                            }
                            uint32_t r[32];
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                float ax = 0.f, ay = 0.f, az = 0.f, aw = 0.f;
                                ax += f0 * a0[i].x; ay += f0 * a0[i].y; az += f0 * a0[i].z; aw += f0 * a0[i].w;
                                ax += f1 * a1[i].x; ay += f1 * a1[i].y; az += f1 * a1[i].z; aw += f1 * a1[i].w;
                                r[4 * i] = __float_as_uint(ax * inv_l); r[4 * i + 1] = __float_as_uint(ay * inv_l);
                                r[4 * i + 2] = __float_as_uint(az * inv_l); r[4 * i + 3] = __float_as_uint(aw * inv_l);
                            }
                            pp_store_rows(out_pair, p.ld_out, warp_row0, N, ch, r, lane);
                        }
                    } else {
                    float M = -INFINITY;
                    for (int c = c_first; c <= c_last; ++c) M = fmaxf(M, pp_ldcg(p.part_ml + (part_slot(c) * PP_SUB + row) * 2));
                    float L = 0.f;
                    for (int c = c_first; c <= c_last; ++c) {
                        const float* ml = p.part_ml + (part_slot(c) * PP_SUB + row) * 2;
                        L += exp2f(pp_ldcg(ml) - M) * pp_ldcg(ml + 1);
                    }
                    const float inv_l = 1.0f / L;
#pragma unroll 1
                    for (int ch = 0; ch < 2; ++ch) {                      // 32 channels at a time, as in the direct epilogue
                        float acc[32];
#pragma unroll
                        for (int i = 0; i < 32; ++i) acc[i] = 0.f;
#pragma unroll 1
                        for (int c = c_first; c <= c_last; ++c) {
                            const size_t ps = part_slot(c);
                            const float* src_o = p.part_o + ps * (PP_SUB * PP_C);
                            float4 v4[8];
#pragma unroll
                            for (int i = 0; i < 8; ++i) v4[i] = pp_ldcg4(src_o + 4 * (size_t)pp_priv_index(w8, ch * 8 + i, lane));     // eight loads in flight
                            const float f = exp2f(pp_ldcg(p.part_ml + (ps * PP_SUB + row) * 2) - M);
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                acc[4 * i] += f * v4[i].x; acc[4 * i + 1] += f * v4[i].y; acc[4 * i + 2] += f * v4[i].z; acc[4 * i + 3] += f * v4[i].w;
                            }
                        }
                        uint32_t r[32];
#pragma unroll
                        for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(acc[i] * inv_l);
                        pp_store_rows(out_pair, p.ld_out, warp_row0, N, ch, r, lane);
                    }
                    }
                }
            }
            if ((warp & 7) == 0) PP_TRACE(240 + min(si, 15), 6 + x);
            if (si == 0) PP_TRACE(224 + warp, 0);
            s = s_next;
            have = have_next;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (CL > 1) cluster_sync_all();            // the peer may still signal this CTA's barriers until it is done too
    if (warp == PP_SOFT_WARPS + 1) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------
// Tuning knobs below are read only by the -DPDSC_TUNING build (common.cuh); the shipped library uses the defaults.
// PDSC_ATTN_CLUSTER: CTAs per cluster, 1 or 2.
static int pp_cluster() { return tuning_int("PDSC_ATTN_CLUSTER", PP_DEFAULT_CLUSTER) == 2 ? 2 : 1; }

static int pp_smem_bytes() { return (int)sizeof(PpSmem) + 1024; }

// clusters of 2 that can be resident at once (one CTA per SM; both CTAs of a cluster sit in one GPC)
#ifndef PDSC_TUNING
static int pp_max_clusters() { return 0; }      // the shipped library never launches the cluster instantiation
#else
static int pp_max_clusters() {
    static thread_local int cached = -1, cached_dev = -1;       // per device, like num_sms()
    int dev = 0;
    cudaGetDevice(&dev);
    if (cached >= 0 && dev == cached_dev) return cached;
    cached_dev = dev;
    opt_in_smem<attn_pp_kernel<2, false>>(pp_smem_bytes());
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * num_sms()); cfg.blockDim = dim3(PP_THREADS); cfg.dynamicSmemBytes = pp_smem_bytes();
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, attn_pp_kernel<2, false>, &cfg) != cudaSuccess || n < 1) { cudaGetLastError(); n = num_sms() / 2; }
    cached = n;
    return n;
}
#endif

// number of work ranges (= CTAs without clusters, = clusters with them)
static int pp_grid(int B, int NU, int NT, int CL) {
    const long long total = (long long)B * NU * NT;
    // PDSC_ATTN_GRID (tuning build): unset = one CTA per SM (stream-K), 0 = one range per work unit (no partials), n = n ranges
    const int knob = tuning_int("PDSC_ATTN_GRID", -1);
    long long g = knob < 0 ? (CL == 1 ? num_sms() : pp_max_clusters()) : (knob == 0 ? (long long)B * NU : knob);
    if (g > total) g = total;
    if (g < 1) g = 1;
    return (int)g;
}

struct PpLayout {
    size_t q, k, v, pts, part_o, part_ml, chain_o, chain_m, counters, n_counters, total;
};
static PpLayout pp_layout(int B, int N) {
    const int NRB = (N + PP_BM - 1) / PP_BM, NT = (N + PP_BN - 1) / PP_BN;
    const int NRBp = (NRB + 1) & ~1;
    const size_t t128 = (size_t)B * 2 * NRBp;
    // partial slots for the larger of the two launch shapes, so the layout does not depend on the cluster knob
    const size_t G = (size_t)std::max(pp_grid(B, NRB, NT, 1), kTuningBuild ? 2 * pp_grid(B, NRBp / 2, NT, 2) : 0);
    PpLayout L;
    size_t off = 0;
    L.q = off; off += t128 * PP_QTILE_BYTES;
    L.k = off; off += t128 * 2 * PP_KTILE_BYTES;
    L.v = off; off += t128 * 2 * PP_VTILE_BYTES;
    L.pts = off; off += align_up(t128 * 2 * PP_PTS_BYTES, 1024);
    L.part_o = off; off += G * 2 * 2 * PP_SUB * PP_C * sizeof(float);
    L.part_ml = off; off += G * 2 * 2 * PP_SUB * 2 * sizeof(float);
    L.chain_o = off; off += (NT > 100 || tuning_int("PDSC_ATTN_CHAIN", -1) >= 0 ? (size_t)PP_CHAIN_SLOTS : 0) * G * 2 * PP_SUB * PP_C * sizeof(float);
    L.chain_m = off; off += (size_t)PP_CHAIN_SLOTS * G * 2 * PP_SUB * sizeof(float);
    L.counters = off; off += align_up((size_t)B * NRBp * 2 * sizeof(int), 1024);
    L.n_counters = (size_t)B * NRBp * 2;
    L.total = off + 1024;
    return L;
}

size_t attn_pp_workspace(int B, int N) { return pp_layout(B, N).total; }
int attn_pp_tiles128(int N) { return 2 * ((((N + PP_BM - 1) / PP_BM) + 1) & ~1); }   // 128-row tiles per pair in the workspace

// counters / n_counters: the stream-K arrival counters of the main kernel.  Whoever writes the operand tiles (qkv_pack_kernel,
// attn_pp_prep_kernel) also zeroes them -- no separate memset on the stream; the main kernel leaves them at zero again.
void attn_pp_tile_pointers(void* ws, int B, int N, uint8_t** Qw, uint8_t** Kw, uint8_t** Vw, float** Pw, int** counters,
                           int* n_counters) {
    const PpLayout L = pp_layout(B, N);
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(ws) + 1023) & ~(uintptr_t)1023);
    *Qw = base + L.q; *Kw = base + L.k; *Vw = base + L.v;
    *Pw = reinterpret_cast<float*>(base + L.pts);
    *counters = reinterpret_cast<int*>(base + L.counters);
    *n_counters = (int)(L.n_counters);
}

// Main kernel on operand tiles already present in the workspace (written by attn_pp_prep_kernel or qkv_pack_kernel).
int launch_attn_pp_main(float* out, int64_t ld_out, int B, int N, int C, int mode, void* ws, size_t ws_bytes, cudaStream_t st) {
    if (C != PP_C) { set_error("nonlocal_attn: C=%d unsupported (kernels are built for C=%d)", C, PP_C); return PDSC_ERR_ARG; }
    const PpLayout L = pp_layout(B, N);
    if (!ws || ws_bytes < L.total) {
        set_error("nonlocal_attn: workspace too small (%zu < %zu)", ws_bytes, L.total);
        return PDSC_ERR_WORKSPACE;
    }
    const int NRB = (N + PP_BM - 1) / PP_BM, NT = (N + PP_BN - 1) / PP_BN;
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(ws) + 1023) & ~(uintptr_t)1023);
    PpParams p;
    p.Qw = base + L.q; p.Kw = base + L.k; p.Vw = base + L.v;
    p.Pw = reinterpret_cast<const float*>(base + L.pts);
    p.out = out; p.ld_out = ld_out;
    p.part_o = reinterpret_cast<float*>(base + L.part_o);
    p.part_ml = reinterpret_cast<float*>(base + L.part_ml);
    p.counters = reinterpret_cast<int*>(base + L.counters);
    p.chain_o = reinterpret_cast<float*>(base + L.chain_o);
    p.chain_m = reinterpret_cast<float*>(base + L.chain_m);
    {   // chain limit (see the work-decomposition comment): on for blocks longer than 100 key tiles (N > 6400);
        // PDSC_ATTN_CHAIN (tuning build) overrides the chain length, 0 = no limit.  At most PP_CHAIN_SLOTS finished
        // chains per segment: very long blocks get proportionally longer chains.
        const int knob = tuning_int("PDSC_ATTN_CHAIN", -1);
        p.chain = knob >= 0 ? knob : (NT > 100 ? 48 : 0);
        if (p.chain > 0 && (NT - 1) / p.chain > PP_CHAIN_SLOTS) p.chain = (NT - 1) / PP_CHAIN_SLOTS + 1;
    }
    const int CL = pp_cluster();
    const int NRBp = (NRB + 1) & ~1;
    p.N = N; p.NT = NT; p.NRBp = NRBp;
    p.NU = CL == 1 ? NRB : NRBp / 2;
    p.total_steps = (long long)B * p.NU * NT;
    // hi/lo product terms per contraction (see the precision note at the top of the file)
    const bool split = mode == PDSC_ATTN_SPLIT16 || mode == PDSC_ATTN_SPLIT16_PV2;
    p.qk_terms = split ? (PP_QK_HH | PP_QK_HL | PP_QK_LH) : PP_QK_HH;
    p.pv_terms = mode == PDSC_ATTN_SPLIT16 ? 3 : (mode == PDSC_ATTN_SPLIT16_PV2 ? 2 : 1);
    if (split) {   // PDSC_ATTN_QK_TERMS (bit mask) / PDSC_ATTN_PV_TERMS (1..3): tuning build only
        const int qk = tuning_int("PDSC_ATTN_QK_TERMS", -1), pv = tuning_int("PDSC_ATTN_PV_TERMS", -1);
        if (qk >= 1 && qk <= 7 && (qk & 1)) p.qk_terms = qk;
        if (pv >= 1 && pv <= 3) p.pv_terms = pv;
    }
    // PDSC_ATTN_DBG (tuning build): selects the debug instantiation; bits 1 / 2 / 4 switch off the compat arithmetic,
    // the exp, the MMAs (timing experiments only, results are then wrong); 8 sets no knob
    // PDSC_ATTN_HINT_SOFT / _ROLE (tuning build): suspend-time hints of the mbarrier waits, ns
    p.hint_soft = (uint32_t)tuning_int("PDSC_ATTN_HINT_SOFT", 20000);
    p.hint_role = (uint32_t)tuning_int("PDSC_ATTN_HINT_ROLE", 20000);
    const int dbg = tuning_int("PDSC_ATTN_DBG", 0);
    p.dbg = dbg;
    if (dbg & 4) p.qk_terms = p.pv_terms = 0;       // no MMAs
    const int smem_bytes = pp_smem_bytes();
    opt_in_smem<attn_pp_kernel<1, false>>(smem_bytes);
    opt_in_smem<attn_pp_kernel<1, false, true>>(smem_bytes);
    opt_in_smem<attn_pp_kernel<1, false, false, true>>(smem_bytes);
    opt_in_smem<attn_pp_kernel<1, false, true, true>>(smem_bytes);
    const bool pv3 = p.pv_terms > 2;
    const int G = pp_grid(B, p.NU, NT, CL);
    p.trace = nullptr;
#ifdef PDSC_TUNING
    opt_in_smem<attn_pp_kernel<2, false>>(smem_bytes);
    opt_in_smem<attn_pp_kernel<2, false, false, true>>(smem_bytes);
    opt_in_smem<attn_pp_kernel<1, true>>(smem_bytes);
    opt_in_smem<attn_pp_kernel<1, true, false, true>>(smem_bytes);
    const char* trace_path = tuning_str("PDSC_ATTN_TRACE");     // debugging aid: timeline of CTA 0, written as text
    if (trace_path) { cudaMalloc(&p.trace, sizeof(long long) * 16 * 256); cudaMemset(p.trace, 0, sizeof(long long) * 16 * 256); }
    if (CL == 1 && (dbg || trace_path)) {
        if (pv3) attn_pp_kernel<1, true, false, true><<<G, PP_THREADS, smem_bytes, st>>>(p);
        else attn_pp_kernel<1, true><<<G, PP_THREADS, smem_bytes, st>>>(p);
    } else if (CL == 2) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(2 * G); cfg.blockDim = dim3(PP_THREADS); cfg.dynamicSmemBytes = smem_bytes; cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        if (pv3) cudaLaunchKernelEx(&cfg, attn_pp_kernel<2, false, false, true>, p);
        else cudaLaunchKernelEx(&cfg, attn_pp_kernel<2, false>, p);
    } else
#endif
    if (p.chain > 0) {
        if (pv3) attn_pp_kernel<1, false, true, true><<<G, PP_THREADS, smem_bytes, st>>>(p);
        else attn_pp_kernel<1, false, true><<<G, PP_THREADS, smem_bytes, st>>>(p);
    } else {
        if (pv3) attn_pp_kernel<1, false, false, true><<<G, PP_THREADS, smem_bytes, st>>>(p);
        else attn_pp_kernel<1, false><<<G, PP_THREADS, smem_bytes, st>>>(p);
    }
#ifdef PDSC_TUNING
    if (trace_path) {
        static long long h[16 * 256];
        cudaDeviceSynchronize();
        cudaMemcpy(h, p.trace, sizeof(h), cudaMemcpyDeviceToHost);
        if (FILE* f = fopen(trace_path, "w")) {
            fprintf(f, "# step kIssued vIssued pfullA pfullB issuedA issuedB compatA compatB sSeenA sSeenB pArrA pArrB vfull kfullNext (clocks relative to step 0 kIssued)\n");
            for (int i = 0; i < 224 && h[i * 16]; ++i) {
                fprintf(f, "%d", i);
                for (int e = 0; e < 14; ++e) fprintf(f, " %lld", h[i * 16 + e] ? h[i * 16 + e] - h[0] : -1);
                fprintf(f, "\n");
            }
            fprintf(f, "# per warp at the first boundary: epilogue done | next segment entered | Q_hi + first compat done | S seen | P handed over | last PV seen | - | last P of segment 0 handed over\n");
            for (int i = 224; i < 240; ++i) {
                fprintf(f, "# warp %d", i - 224);
                for (int e = 0; e < 8; ++e) fprintf(f, " %lld", h[i * 16 + e] ? h[i * 16 + e] - h[0] : -1);
                fprintf(f, "\n");
            }
            fprintf(f, "# segment startA startB bodyDoneA bodyDoneB lastPvA lastPvB epiDoneA epiDoneB\n");
            for (int i = 240; i < 256 && h[i * 16]; ++i) {
                fprintf(f, "# seg %d", i - 240);
                for (int e = 0; e < 8; ++e) fprintf(f, " %lld", h[i * 16 + e] ? h[i * 16 + e] - h[0] : -1);
                fprintf(f, "\n");
            }
            fclose(f);
        }
        cudaFree(p.trace);
    }
#endif
    count_launches(1);
    return check_launch("nonlocal_attn(tcgen05 ping-pong)");
}

int launch_attn_pp(const float* q, const float* k, const float* v, int64_t ld, const float* src, const float* tgt,
                   float* out, int64_t ld_out, int B, int N, int C, float sigma_spat, int mode, void* ws,
                   size_t ws_bytes, cudaStream_t st) {
    if (C != PP_C) { set_error("nonlocal_attn: C=%d unsupported (kernels are built for C=%d)", C, PP_C); return PDSC_ERR_ARG; }
    const PpLayout L = pp_layout(B, N);
    if (!ws || ws_bytes < L.total) {
        set_error("nonlocal_attn: workspace too small (%zu < %zu)", ws_bytes, L.total);
        return PDSC_ERR_WORKSPACE;
    }
    const int NT128 = attn_pp_tiles128(N);
    uint8_t *Qw, *Kw, *Vw;
    float* Pw;
    int* counters;
    int n_counters;
    attn_pp_tile_pointers(ws, B, N, &Qw, &Kw, &Vw, &Pw, &counters, &n_counters);
    opt_in_smem<attn_pp_prep_kernel>(128 * 129 * 4);
    const float qscale = 1.4426950408889634f / sqrtf((float)C);
    attn_pp_prep_kernel<<<dim3(NT128, B), 256, 128 * 129 * 4, st>>>(q, k, v, ld, src, tgt, N, NT128, qscale, 1.0f / sigma_spat,
                                                                      Qw, Kw, Vw, Pw, counters, n_counters);
    count_launches(1);
    return launch_attn_pp_main(out, ld_out, B, N, C, mode, ws, ws_bytes, st);
}

}  // namespace pdsc
