// (a) Compatibility-weighted non-local attention on the 5th-generation tensor cores.
// Replaces models/PointDSC.py:150-153 + NonLocalBlock.forward :39-42.
//
//   S = Q K^T / sqrt(C)            tcgen05.mma kind::f16; Q lives in TMEM (A operand), K tiles are
//                                  bulk-async (TMA engine) staged, 128B-swizzled, double buffered
//   W = softmax_j(compat_ij S_ij)  CUDA cores: compat recomputed from 6 floats per correspondence
//                                  (never stored), online softmax with a lazily raised reference max
//   msg = W V                      tcgen05.mma with P read from TMEM (aliasing S) and V^T from smem
//
// Precision (PDSC_ATTN_SPLIT16): every fp32 operand x is split into fp16 hi + fp16 lo (22 mantissa
// bits) and each contraction is issued as hi*hi + hi*lo + lo*hi with fp32 accumulation, which keeps
// the 12-layer encoder inside the 1e-4 confidence tolerance (single-pass bf16/tf32/fp16 do not).
// PDSC_ATTN_FP16 issues only hi*hi.
//
// One CTA = 128 query rows of one pair, looping over 128-key tiles:
//   warp 0      producer: K tiles (2 stages) and key points (2 stages)
//   warp 1      MMA issuer (one elected lane), owns the TMEM allocation
//   warp 2      producer of the V^T tiles (hi and lo halves on separate barriers)
//   warps 3..   NWG softmax warpgroups; thread <-> query row, warpgroup g takes keys [g*128/NWG, ...)
// TMEM columns: S0 [0,128) S1 [128,256) O [256,384) Q [384,512) (fp16 pairs: hi 64 cols, lo 64 cols).
// All warpgroups accumulate into the one O, so they share the reference max of a row: the row maxima
// of a tile are exchanged through shared memory and an mbarrier before P is formed.
#include "common.cuh"
#include "sm100.cuh"

#include <math.h>
#include <stdlib.h>

namespace pdsc {

using namespace sm100;

constexpr int TC_C = 128;            // channels (head dim); kernels are built for C = 128
constexpr int TC_BM = 128;           // query rows per CTA
constexpr int TC_BN = 128;           // keys per tile
constexpr int TC_MAXWG = 4;
constexpr uint32_t TC_BLOCK_BYTES = 128 * 128;            // one swizzled block: 128 rows x 64 fp16
constexpr uint32_t TC_TILE_BYTES = 4 * TC_BLOCK_BYTES;    // [hi|lo][k-chunk 0|1]
constexpr uint32_t TC_PTS_FLOATS = (TC_BN / 2) * 12;      // per key PAIR: sx0 sx1 sy0 sy1 sz0 sz1 tx0 tx1 ty0 ty1 tz0 tz1
constexpr uint32_t TC_PTS_BYTES = TC_PTS_FLOATS * 4;
constexpr float TC_RESCALE_LOG2 = 8.0f;                   // lazy rescale threshold (log2 units)
constexpr uint32_t TM_S = 0, TM_O = 256, TM_Q = 384;

struct TcSmem {
    uint8_t k[2][TC_TILE_BYTES];
    uint8_t v[TC_TILE_BYTES];
    float4 pts[2][TC_BN / 2][3];
    float xch[2][TC_MAXWG][TC_BM];
    float lsum[TC_MAXWG][TC_BM];
    uint64_t k_full[2], k_empty[2], vh_full, vl_full, vh_empty, vl_empty;
    uint64_t pts_full[2], pts_empty[2], s_full[2], p_full[2], xch_bar[2], pv_done, q_ready;
    uint32_t tmem_base;
};

// byte offset of (row r, 16-byte chunk c) inside a 128-row x 128-byte block with the 128B swizzle
__host__ __device__ inline uint32_t swz(uint32_t r, uint32_t c) { return r * 128u + ((c ^ (r & 7u)) << 4); }

// ------------------------------------------------------------------------------------------------
// Prep: fp32 q,k,v [B,N,C] -> fp16 hi/lo tiles.  K and V^T tiles are written as the exact
// shared-memory image the MMAs read (so a tile is ONE contiguous bulk copy; V transposed so that
// keys become the contraction dim).  Q tiles are row-linear [row][hi 64 words | lo 64 words] (they
// go to TMEM through registers).  q is pre-multiplied by log2(e)/sqrt(C); the key points are
// divided by sigma_d, so that compat = sat(1 - (|ds| - |dt|)^2), and packed per key pair.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void split8(const float (&x)[8], uint4& hi, uint4& lo) {
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

__global__ void __launch_bounds__(256) attn_prep_kernel(const float* __restrict__ q, const float* __restrict__ k,
                                                        const float* __restrict__ v, int64_t ld,
                                                        const float* __restrict__ src, const float* __restrict__ tgt,
                                                        int N, int NT, float qscale, float inv_sigma,
                                                        uint8_t* __restrict__ Qw, uint8_t* __restrict__ Kw,
                                                        uint8_t* __restrict__ Vw, float* __restrict__ Pw) {
    extern __shared__ __align__(16) float vt[];   // [128 keys][129] staging for the V transpose
    const int t = blockIdx.x, b = blockIdx.y;
    const int n0 = t * TC_BN;
    const size_t tile = (size_t)b * NT + t;
    uint8_t* qd = Qw + tile * TC_TILE_BYTES;
    uint8_t* kd = Kw + tile * TC_TILE_BYTES;
    uint8_t* vd = Vw + tile * TC_TILE_BYTES;
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
        split8(xq, hi, lo);
        *reinterpret_cast<uint4*>(qd + r * 512 + c8 * 16) = hi;
        *reinterpret_cast<uint4*>(qd + r * 512 + 256 + c8 * 16) = lo;
        const uint32_t off = (uint32_t)(c8 >> 3) * TC_BLOCK_BYTES + swz(r, c8 & 7);
        split8(xk, hi, lo);
        *reinterpret_cast<uint4*>(kd + off) = hi;
        *reinterpret_cast<uint4*>(kd + 2 * TC_BLOCK_BYTES + off) = lo;
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
        split8(x, hi, lo);
        const uint32_t off = (uint32_t)(k8 >> 3) * TC_BLOCK_BYTES + swz(ch, k8 & 7);
        *reinterpret_cast<uint4*>(vd + off) = hi;
        *reinterpret_cast<uint4*>(vd + 2 * TC_BLOCK_BYTES + off) = lo;
    }
    if (threadIdx.x < TC_BN) {
        // packed-pair SoA layout so the softmax loop can use FADD2/FFMA2 on two keys at once
        const int n = min(n0 + (int)threadIdx.x, N - 1);
        const float* s = src + ((size_t)b * N + n) * 3;
        const float* g = tgt + ((size_t)b * N + n) * 3;
        float* d = Pw + tile * TC_PTS_FLOATS + (threadIdx.x >> 1) * 12 + (threadIdx.x & 1);
        d[0] = s[0] * inv_sigma; d[2] = s[1] * inv_sigma; d[4] = s[2] * inv_sigma;
        d[6] = g[0] * inv_sigma; d[8] = g[1] * inv_sigma; d[10] = g[2] * inv_sigma;
    }
}

// ------------------------------------------------------------------------------------------------
// MMA issue helpers (one thread).  A operands come from TMEM (fp16 pairs, 8 columns per K=16 step),
// B operands from shared memory as two 64-element swizzled blocks per 128-element contraction.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t b_off(int kk) { return (uint32_t)(kk >> 2) * TC_BLOCK_BYTES + (uint32_t)(kk & 3) * 32u; }

// S[128 x 128 keys] = Q K^T: `terms` = 3 (hi*hi + hi*lo + lo*hi) or 1 (hi*hi)
__device__ __forceinline__ void issue_qk(uint32_t s_tmem, uint32_t q_tmem, uint32_t k_base, int terms) {
    if (terms <= 0) return;
    constexpr uint32_t idesc = umma_idesc_f16(128, 128);
    uint32_t acc = 0;
    for (int term = 0; term < terms; ++term) {
        const uint32_t a_part = (term == 2) ? 64u : 0u;                   // lo of Q in the third term
        const uint32_t b_part = (term == 1) ? 2u * TC_BLOCK_BYTES : 0u;   // lo of K in the second term
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
            umma_ts(s_tmem, q_tmem + a_part + (uint32_t)kk * 8u, umma_desc_sw128_kmajor(k_base + b_part + b_off(kk)), idesc, acc);
            acc = 1;
        }
    }
}
// O[128 x 128 ch] (+)= P[128 x 128 keys] V: the two terms that read the hi half of V^T ...
__device__ __forceinline__ void issue_pv_vhi(uint32_t o_tmem, uint32_t p_tmem, uint32_t v_base, int terms, uint32_t acc) {
    if (terms <= 0) return;
    constexpr uint32_t idesc = umma_idesc_f16(128, 128);
    for (int term = 0; term < (terms > 1 ? 2 : 1); ++term) {
        const uint32_t a_part = term ? 64u : 0u;                          // P_lo in the second one
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
            umma_ts(o_tmem, p_tmem + a_part + (uint32_t)kk * 8u, umma_desc_sw128_kmajor(v_base + b_off(kk)), idesc, acc);
            acc = 1;
        }
    }
}
// ... and the one that reads the lo half (P_hi * V_lo)
__device__ __forceinline__ void issue_pv_vlo(uint32_t o_tmem, uint32_t p_tmem, uint32_t v_base) {
    constexpr uint32_t idesc = umma_idesc_f16(128, 128);
#pragma unroll
    for (int kk = 0; kk < 8; ++kk)
        umma_ts(o_tmem, p_tmem + (uint32_t)kk * 8u, umma_desc_sw128_kmajor(v_base + 2u * TC_BLOCK_BYTES + b_off(kk)), idesc, 1u);
}

__device__ __forceinline__ float fma_sat(float a, float b, float c) {
    float d;
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

// compat of this thread's query point against NP key pairs -> w[NP] (two keys per float2)
template <int NP>
__device__ __forceinline__ void compat_pairs(const float4 (*pts)[3], const float2 (&qp)[6], float2 (&w)[NP]) {
#pragma unroll
    for (int i = 0; i < NP; ++i) {
        const float4 pa = pts[i][0], pb = pts[i][1], pc = pts[i][2];
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
        w[i] = make_float2(fma_sat(-c.x, c.x, 1.0f), fma_sat(-c.y, c.y, 1.0f));   // max(0, 1 - c^2)
    }
}

// ------------------------------------------------------------------------------------------------
// Main kernel.
// ------------------------------------------------------------------------------------------------
template <int NWG>
__global__ void __launch_bounds__(96 + 128 * NWG, 1) attn_tc_kernel(const uint8_t* __restrict__ Qw,
                                                                    const uint8_t* __restrict__ Kw,
                                                                    const uint8_t* __restrict__ Vw,
                                                                    const float* __restrict__ Pw, float* __restrict__ out,
                                                                    int64_t ld_out, int N, int NT, int terms, int dbg) {
    constexpr int KPW = TC_BN / NWG;         // keys per warpgroup per tile
    constexpr int NP = KPW / 2;              // key pairs per thread per tile
    constexpr int CPW = TC_C / NWG;          // O columns / output channels / Q words per warpgroup
    constexpr uint32_t NSOFT_WARPS = 4 * NWG;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // the 128B swizzle is anchored at 1024-byte boundaries of the shared window
    TcSmem& sm = *reinterpret_cast<TcSmem*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
    const int qt = blockIdx.x, b = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t tile0 = (size_t)b * NT;

    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(&sm.k_full[i], 1); mbar_init(&sm.k_empty[i], 1);
            mbar_init(&sm.pts_full[i], 1); mbar_init(&sm.pts_empty[i], NSOFT_WARPS);
            mbar_init(&sm.s_full[i], 1); mbar_init(&sm.p_full[i], NSOFT_WARPS);
            mbar_init(&sm.xch_bar[i], NSOFT_WARPS);
        }
        mbar_init(&sm.vh_full, 1); mbar_init(&sm.vl_full, 1); mbar_init(&sm.vh_empty, 1); mbar_init(&sm.vl_empty, 1);
        mbar_init(&sm.pv_done, 1); mbar_init(&sm.q_ready, NSOFT_WARPS);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(&sm.tmem_base, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = sm.tmem_base;

    if (warp == 0) {
        // ============================== producer: K tiles, key points ==============================
        if (lane == 0) {
            for (int j = 0; j < NT; ++j) {
                const int st = j & 1;
                const uint32_t par = ((j >> 1) - 1) & 1;                 // previous use of this stage
                if (j >= 2) mbar_wait_role(&sm.k_empty[st], par);        // QK(j-2) has consumed it
                mbar_arrive_expect_tx(&sm.k_full[st], TC_TILE_BYTES);
                const uint8_t* g = Kw + (tile0 + j) * TC_TILE_BYTES;
                bulk_g2s(sm.k[st], g, TC_TILE_BYTES / 2, &sm.k_full[st]);
                bulk_g2s(sm.k[st] + TC_TILE_BYTES / 2, g + TC_TILE_BYTES / 2, TC_TILE_BYTES / 2, &sm.k_full[st]);
                if (j >= 2) mbar_wait_role(&sm.pts_empty[st], par);      // compat(j-2) has been computed
                mbar_arrive_expect_tx(&sm.pts_full[st], TC_PTS_BYTES);
                bulk_g2s(sm.pts[st], Pw + (tile0 + j) * TC_PTS_FLOATS, TC_PTS_BYTES, &sm.pts_full[st]);
            }
        }
    } else if (warp == 2) {
        // ============================== producer: V^T tiles, hi and lo halves ======================
        if (lane == 0) {
            for (int j = 0; j < NT; ++j) {
                const uint8_t* g = Vw + (tile0 + j) * TC_TILE_BYTES;
                if (j > 0) mbar_wait_role(&sm.vh_empty, (j - 1) & 1);
                mbar_arrive_expect_tx(&sm.vh_full, TC_TILE_BYTES / 2);
                bulk_g2s(sm.v, g, TC_TILE_BYTES / 2, &sm.vh_full);
                if (terms > 1) {
                    if (j > 0) mbar_wait_role(&sm.vl_empty, (j - 1) & 1);
                    mbar_arrive_expect_tx(&sm.vl_full, TC_TILE_BYTES / 2);
                    bulk_g2s(sm.v + TC_TILE_BYTES / 2, g + TC_TILE_BYTES / 2, TC_TILE_BYTES / 2, &sm.vl_full);
                }
            }
        }
    } else if (warp == 1) {
        // ============================== MMA issuer ==============================
        if (lane == 0) {
            const uint32_t k_addr[2] = {smem_u32(sm.k[0]), smem_u32(sm.k[1])};
            const uint32_t v_addr = smem_u32(sm.v);
            mbar_wait_role(&sm.q_ready, 0);
            mbar_wait_role(&sm.k_full[0], 0);
            tc_fence_after();
            issue_qk(tmem + TM_S, tmem + TM_Q, k_addr[0], terms);
            umma_commit(&sm.s_full[0]);
            umma_commit(&sm.k_empty[0]);
            for (int j = 0; j < NT; ++j) {
                const int buf = j & 1;
                if (j + 1 < NT) {
                    mbar_wait_role(&sm.k_full[buf ^ 1], ((j + 1) >> 1) & 1);
                    tc_fence_after();
                    issue_qk(tmem + TM_S + (uint32_t)(buf ^ 1) * 128u, tmem + TM_Q, k_addr[buf ^ 1], terms);
                    umma_commit(&sm.s_full[buf ^ 1]);
                    umma_commit(&sm.k_empty[buf ^ 1]);
                }
                mbar_wait_role(&sm.vh_full, j & 1);
                mbar_wait_role(&sm.p_full[buf], (j >> 1) & 1);
                tc_fence_after();
                issue_pv_vhi(tmem + TM_O, tmem + TM_S + (uint32_t)buf * 128u, v_addr, terms, j > 0 ? 1u : 0u);
                umma_commit(&sm.vh_empty);
                if (terms > 1) {
                    mbar_wait_role(&sm.vl_full, j & 1);
                    tc_fence_after();
                    issue_pv_vlo(tmem + TM_O, tmem + TM_S + (uint32_t)buf * 128u, v_addr);
                    umma_commit(&sm.vl_empty);
                }
                umma_commit(&sm.pv_done);
            }
        }
    } else {
        // ============================== softmax warpgroups ==============================
        const int wg = (warp - 3) >> 2;              // keys [KPW wg, KPW wg + KPW) of every tile
        const int quarter = warp & 3;                // TMEM lane quarter this warp may access
        const int row = quarter * 32 + lane;         // query row inside the CTA tile
        const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
        const int grow = qt * TC_BM + row;
        const int qrow = min(grow, N - 1);

        // ---- Q tile -> TMEM (this warpgroup's CPW words of every row) ----
        {
            const uint4* qsrc = reinterpret_cast<const uint4*>(Qw + (tile0 + qt) * TC_TILE_BYTES + (size_t)row * 512 + wg * CPW * 4);
#pragma unroll
            for (int h = 0; h < CPW / 32; ++h) {
                uint32_t r[32];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const uint4 x = __ldg(qsrc + h * 8 + i);
                    r[4 * i] = x.x; r[4 * i + 1] = x.y; r[4 * i + 2] = x.z; r[4 * i + 3] = x.w;
                }
                tmem_st32(tmem + lane_base + TM_Q + (uint32_t)(wg * CPW + h * 32), r);
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.q_ready);
        }
        float2 qp[6];
        {
            const float* p = Pw + (tile0 + (qrow >> 7)) * TC_PTS_FLOATS + ((qrow & 127) >> 1) * 12 + (qrow & 1);
#pragma unroll
            for (int i = 0; i < 6; ++i) qp[i] = make_float2(p[2 * i], p[2 * i]);
        }
        const int kp0 = wg * NP;                     // first key pair of this warpgroup inside a tile
        const uint32_t o_own = tmem + lane_base + TM_O + (uint32_t)(wg * CPW);
        float m = 0.f, l = 0.f;                      // reference max (log2 units; compat = 0 keys sit at 0) and row sum
        for (int j = 0; j < NT; ++j) {
            const int buf = j & 1;
            const uint32_t par = (j >> 1) & 1;
            // compatibilities of this thread's row against the tile's keys (needs only the points)
            float2 e[NP];
            mbar_wait(&sm.pts_full[buf], par);
            if (dbg & 1) {
#pragma unroll
                for (int i = 0; i < NP; ++i) e[i] = make_float2(0.5f, 0.25f);
            } else {
                compat_pairs<NP>(&sm.pts[buf][kp0], qp, e);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.pts_empty[buf]);
            mbar_wait(&sm.s_full[buf], par);
            tc_fence_after();
            const uint32_t s_buf = tmem + lane_base + TM_S + (uint32_t)buf * 128u;
            // e = compat * S - m  (log2 units), and its maximum over this thread's keys
            float emax = -INFINITY;
            const float2 neg_m = make_float2(-m, -m);
#pragma unroll
            for (int h = 0; h < KPW / 32; ++h) {
                uint32_t r[32];
                tmem_ld32(s_buf + (uint32_t)(wg * KPW + h * 32), r);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float2 ev = __ffma2_rn(e[h * 16 + i], make_float2(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1])), neg_m);
                    e[h * 16 + i] = ev;
                    emax = fmaxf(emax, fmaxf(ev.x, ev.y));
                }
            }
            const int nvalid = N - j * TC_BN;        // keys of this tile that exist
            if (nvalid < TC_BN) {                    // last, partial tile: keys past N do not exist
                emax = -INFINITY;
#pragma unroll
                for (int i = 0; i < NP; ++i) {
                    const int key = wg * KPW + 2 * i;
                    if (key >= nvalid) e[i].x = -INFINITY;
                    if (key + 1 >= nvalid) e[i].y = -INFINITY;
                    emax = fmaxf(emax, fmaxf(e[i].x, e[i].y));
                }
            }
            // all warpgroups must use the same reference max: exchange the row maxima
            sm.xch[buf][wg][row] = emax;
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.xch_bar[buf]);
            mbar_wait(&sm.xch_bar[buf], par);
#pragma unroll
            for (int g = 0; g < NWG; ++g) emax = fmaxf(emax, sm.xch[buf][g][row]);
            // Lazy reference max: raised only when the row maximum exceeds it by more than 2^8; on
            // the first tile it is also lowered for rows whose logits are all far below zero.
            const bool need = (emax > TC_RESCALE_LOG2) || (j == 0 && emax < -TC_RESCALE_LOG2 && emax > -INFINITY);
            if (__any_sync(0xffffffffu, need)) {
                const float shift = need ? emax : 0.f;
                const float f = fast_exp2(-shift);
                m += shift;
                l *= f;
                const float2 ns = make_float2(-shift, -shift);
#pragma unroll
                for (int i = 0; i < NP; ++i) e[i] = __fadd2_rn(e[i], ns);
                if (j > 0) {
                    mbar_wait(&sm.pv_done, (j - 1) & 1);      // PV(j-1) has finished accumulating into O
                    tc_fence_after();
#pragma unroll
                    for (int h = 0; h < CPW / 32; ++h) {
                        uint32_t r[32];
                        tmem_ld32(o_own + h * 32, r);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * f);
                        tmem_st32(o_own + h * 32, r);
                    }
                }
            }
            // P = exp2(e), split into fp16 hi/lo, packed two keys per TMEM column:
            // hi pairs at S columns [0,64), lo pairs at [64,128) of the same buffer
            float2 psum = make_float2(0.f, 0.f);
#pragma unroll
            for (int c = 0; c < NP / 16; ++c) {
                uint32_t hi[16], lo[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float2 ev = e[c * 16 + i];
                    const float2 p = (dbg & 2) ? make_float2(1.0f + 1e-3f * ev.x, 1.0f + 1e-3f * ev.y) : make_float2(fast_exp2(ev.x), fast_exp2(ev.y));
                    psum = __fadd2_rn(psum, p);
                    hi[i] = pack_half2(p.x, p.y);
                    const float2 back = unpack_half2(hi[i]);
                    const float2 rem = __fadd2_rn(p, make_float2(-back.x, -back.y));
                    lo[i] = pack_half2(rem.x, rem.y);
                }
                tmem_st16(s_buf + (uint32_t)(kp0 + c * 16), hi);
                if (terms > 1) tmem_st16(s_buf + 64u + (uint32_t)(kp0 + c * 16), lo);
            }
            l += psum.x + psum.y;
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.p_full[buf]);
        }
        // ---- epilogue: msg = O / L, this warpgroup writes channels [CPW wg, CPW wg + CPW) ----
        sm.lsum[wg][row] = l;
        named_bar_sync(1, 128 * NWG);
        float L = 0.f;
#pragma unroll
        for (int g = 0; g < NWG; ++g) L += sm.lsum[g][row];
        const float inv_l = 1.0f / L;
        mbar_wait(&sm.pv_done, (NT - 1) & 1);
        tc_fence_after();
        float* orow = out + ((size_t)b * N + qrow) * ld_out + wg * CPW;
#pragma unroll
        for (int h = 0; h < CPW / 32; ++h) {
            uint32_t r[32];
            tmem_ld32(o_own + h * 32, r);
            tmem_ld_wait();
            if (grow < N) {
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    float4 o;
                    o.x = __uint_as_float(r[i]) * inv_l;
                    o.y = __uint_as_float(r[i + 1]) * inv_l;
                    o.z = __uint_as_float(r[i + 2]) * inv_l;
                    o.w = __uint_as_float(r[i + 3]) * inv_l;
                    *reinterpret_cast<float4*>(orow + h * 32 + i) = o;
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------------
static size_t tc_tiles(int N) { return (size_t)((N + TC_BN - 1) / TC_BN); }

size_t attn_workspace(int B, int N, int C, int mode) {
    (void)C;
    if (mode == PDSC_ATTN_FP32) return 0;
    const size_t tiles = (size_t)B * tc_tiles(N);
    return tiles * (3 * (size_t)TC_TILE_BYTES + TC_PTS_BYTES) + 1024;
}

static int softmax_warpgroups() {
    static int nwg = 0;
    if (!nwg) {
        const char* e = getenv("PDSC_ATTN_NWG");     // tuning knob: 2 or 4 softmax warpgroups
        nwg = (e && atoi(e) == 2) ? 2 : 4;
    }
    return nwg;
}

int launch_attn_tc(const float* q, const float* k, const float* v, int64_t ld, const float* src, const float* tgt,
                   float* out, int64_t ld_out, int B, int N, int C, float sigma_spat, int mode, void* ws,
                   size_t ws_bytes, cudaStream_t st) {
    if (C != TC_C) { set_error("nonlocal_attn: C=%d unsupported (kernels are built for C=%d)", C, TC_C); return PDSC_ERR_ARG; }
    if (!ws || ws_bytes < attn_workspace(B, N, C, mode)) {
        set_error("nonlocal_attn: workspace too small (%zu < %zu)", ws_bytes, attn_workspace(B, N, C, mode));
        return PDSC_ERR_WORKSPACE;
    }
    const int NT = (int)tc_tiles(N);
    const size_t tiles = (size_t)B * NT;
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(ws) + 1023) & ~(uintptr_t)1023);
    uint8_t* Qw = base;
    uint8_t* Kw = Qw + tiles * TC_TILE_BYTES;
    uint8_t* Vw = Kw + tiles * TC_TILE_BYTES;
    float* Pw = reinterpret_cast<float*>(Vw + tiles * TC_TILE_BYTES);
    const int smem_bytes = (int)sizeof(TcSmem) + 1024;
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(attn_prep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 129 * 4);
        cudaFuncSetAttribute(attn_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
        cudaFuncSetAttribute(attn_tc_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
        attr_set = true;
    }
    const float qscale = 1.4426950408889634f / sqrtf((float)C);
    attn_prep_kernel<<<dim3(NT, B), 256, 128 * 129 * 4, st>>>(q, k, v, ld, src, tgt, N, NT, qscale, 1.0f / sigma_spat, Qw, Kw,
                                                              Vw, Pw);
    int terms = mode == PDSC_ATTN_SPLIT16 ? 3 : 1;
    static int dbg = -1;                            // PDSC_ATTN_DBG: timing experiments only (results are wrong)
    if (dbg < 0) { const char* e = getenv("PDSC_ATTN_DBG"); dbg = e ? atoi(e) : 0; }
    if (dbg & 4) terms = 0;                         // no MMAs
    if (softmax_warpgroups() == 2)
        attn_tc_kernel<2><<<dim3(NT, B), 96 + 256, smem_bytes, st>>>(Qw, Kw, Vw, Pw, out, ld_out, N, NT, terms, dbg);
    else
        attn_tc_kernel<4><<<dim3(NT, B), 96 + 512, smem_bytes, st>>>(Qw, Kw, Vw, Pw, out, ld_out, N, NT, terms, dbg);
    count_launches(2);
    return check_launch("nonlocal_attn(tcgen05)");
}

}  // namespace pdsc
