// (a) Compatibility-weighted non-local attention on the 5th-generation tensor cores.
// Replaces models/PointDSC.py:150-153 + NonLocalBlock.forward :39-42.
//
//   S = Q K^T / sqrt(C)            tcgen05.mma kind::f16, operands in shared memory (bulk-async/TMA
//                                  staged, 128B-swizzled K-major tiles), fp32 accumulators in TMEM
//   W = softmax_j(compat_ij S_ij)  CUDA cores: compat recomputed from 6 floats per correspondence,
//                                  online softmax with lazy rescale, one query row per thread
//   msg = W V                      tcgen05.mma with P read from TMEM (aliasing S) and V^T from smem
//
// Precision (PDSC_ATTN_SPLIT16): every fp32 operand x is split into fp16 hi + fp16 lo (22 mantissa
// bits) and each contraction is issued as hi*hi + hi*lo + lo*hi with fp32 accumulation, which keeps
// the 12-layer encoder inside the 1e-4 confidence tolerance (single-pass bf16/tf32/fp16 do not).
// PDSC_ATTN_FP16 issues only hi*hi.
//
// One CTA = 128 query rows of one pair, looping over 128-key tiles:
//   warp 0      producer: cp.async.bulk of the pre-swizzled Q / K tiles and key points
//   warp 1      MMA issuer (one elected lane), owns the TMEM allocation
//   warp 2      producer of the V^T tiles
//   warps 3..   NWG softmax warpgroups; thread <-> query row, warpgroup g takes keys [g*128/NWG, ...)
// S is double buffered in TMEM so S(j+1)=QK(j+1)^T overlaps softmax(j); PV(j) follows softmax(j).
#include "common.cuh"
#include "sm100.cuh"

#include <math.h>
#include <stdlib.h>

namespace pdsc {

using namespace sm100;

constexpr int TC_C = 128;            // channels (head dim); kernels are built for C = 128
constexpr int TC_BM = 128;           // query rows per CTA
constexpr int TC_BN = 128;           // keys per tile
constexpr int TC_THREADS = 96 + 2 * 128;   // K producer, MMA issuer, V producer, 2 softmax warpgroups
constexpr uint32_t TC_BLOCK_BYTES = 128 * 128;            // one swizzled block: 128 rows x 64 fp16
constexpr uint32_t TC_TILE_BYTES = 4 * TC_BLOCK_BYTES;    // [hi|lo][k-chunk 0|1]
constexpr uint32_t TC_PTS_FLOATS = (TC_BN / 2) * 12;      // per key PAIR: sx0 sx1 sy0 sy1 sz0 sz1 tx0 tx1 ty0 ty1 tz0 tz1
constexpr uint32_t TC_PTS_BYTES = TC_PTS_FLOATS * 4;
constexpr float TC_RESCALE_LOG2 = 8.0f;                   // lazy rescale threshold (log2 units)

struct TcSmem {
    uint8_t q[TC_TILE_BYTES];
    uint8_t k[TC_TILE_BYTES];
    uint8_t v[TC_TILE_BYTES];
    float4 pts[2][TC_BN / 2][3];
    float mfin[2][TC_BM];
    float lsum[2][TC_BM];
    uint64_t q_full, k_full, k_empty, v_full, v_empty;
    uint64_t s_full[2], pts_full[2], pv_done[2], p_full[2][2];
    uint32_t tmem_base;
};

// byte offset of (row r, 16-byte chunk c) inside a 128-row x 128-byte block with the 128B swizzle
__host__ __device__ inline uint32_t swz(uint32_t r, uint32_t c) { return r * 128u + ((c ^ (r & 7u)) << 4); }

// ------------------------------------------------------------------------------------------------
// Prep: fp32 q,k,v [B,N,C] -> fp16 hi/lo tiles in the exact shared-memory image the MMAs read
// (so a tile is ONE contiguous bulk copy), V transposed (keys become the contraction dim), and
// the keypoints packed 8 floats per correspondence.  q is pre-multiplied by log2(e)/sqrt(C).
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
                                                        int N, int NT, float qscale, uint8_t* __restrict__ Qw,
                                                        uint8_t* __restrict__ Kw, uint8_t* __restrict__ Vw,
                                                        float* __restrict__ Pw) {
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
        const uint32_t off = (uint32_t)(c8 >> 3) * TC_BLOCK_BYTES + swz(r, c8 & 7);
        split8(xq, hi, lo);
        *reinterpret_cast<uint4*>(qd + off) = hi;
        *reinterpret_cast<uint4*>(qd + 2 * TC_BLOCK_BYTES + off) = lo;
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
        d[0] = s[0]; d[2] = s[1]; d[4] = s[2]; d[6] = g[0]; d[8] = g[1]; d[10] = g[2];
    }
}

// One 64 KB tile as several bulk copies in flight at once (a single copy is served serially).
#ifndef TC_COPY_SPLIT
#define TC_COPY_SPLIT 1
#endif
__device__ __forceinline__ void bulk_tile(uint8_t* dst, const uint8_t* src, uint64_t* bar) {
    constexpr uint32_t part = TC_TILE_BYTES / TC_COPY_SPLIT;
#pragma unroll
    for (int i = 0; i < TC_COPY_SPLIT; ++i) bulk_g2s(dst + i * part, src + i * part, part, bar);
}

// ------------------------------------------------------------------------------------------------
// MMA issue helpers (one thread)
// ------------------------------------------------------------------------------------------------
// D[128x128] (+)= A[128 x 128ch] * B[128 x 128ch]^T, both operands in shared memory as two
// 64-channel swizzled blocks; `terms` = 3 (hi*hi + hi*lo + lo*hi) or 1 (hi*hi).
__device__ __forceinline__ void issue_qk(uint32_t d_tmem, uint32_t a_base, uint32_t b_base, int terms) {
    constexpr uint32_t idesc = umma_idesc_f16(128, 128);
    uint32_t acc = 0;
    for (int term = 0; term < terms; ++term) {
        const uint32_t a_part = (term == 2) ? 2u * TC_BLOCK_BYTES : 0u;   // lo of A in the third term
        const uint32_t b_part = (term == 1) ? 2u * TC_BLOCK_BYTES : 0u;   // lo of B in the second term
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
            const uint32_t koff = (uint32_t)(kk >> 2) * TC_BLOCK_BYTES + (uint32_t)(kk & 3) * 32u;
            umma_ss(d_tmem, umma_desc_sw128_kmajor(a_base + a_part + koff), umma_desc_sw128_kmajor(b_base + b_part + koff),
                    idesc, acc);
            acc = 1;
        }
    }
}
// O_g[128x128ch] (+)= P_g[128 x 64 keys] * Vt[128ch x 64 keys]^T for the 64 keys of warpgroup g.
// P comes from TMEM (packed fp16: hi at p_tmem, lo 32 columns further, inside the warpgroup's own
// 64-column half of the S buffer), V^T from the g-th 64-key block in shared memory.
__device__ __forceinline__ void issue_pv_half(uint32_t o_tmem, uint32_t p_tmem, uint32_t v_block, int terms, uint32_t acc) {
    constexpr uint32_t idesc = umma_idesc_f16(128, 128);
    for (int term = 0; term < terms; ++term) {
        const uint32_t p_part = (term == 2) ? 32u : 0u;
        const uint32_t v_part = (term == 1) ? 2u * TC_BLOCK_BYTES : 0u;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            umma_ts(o_tmem, p_tmem + (uint32_t)kk * 8u + p_part, umma_desc_sw128_kmajor(v_block + v_part + (uint32_t)kk * 32u),
                    idesc, acc);
            acc = 1;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Main kernel.  TMEM columns: S0 [0,128) S1 [128,256) O_0 [256,384) O_1 [384,512).
// The two softmax warpgroups are fully independent: warpgroup g owns keys [64g,64g+64) of every
// tile, its own running max / sum and its own accumulator O_g; they only meet in the epilogue
// (msg = (O_0 2^(m_0-M) + O_1 2^(m_1-M)) / L).  No per-tile exchange, no CTA-wide barrier.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TC_THREADS, 1) attn_tc_kernel(const uint8_t* __restrict__ Qw,
                                                                const uint8_t* __restrict__ Kw,
                                                                const uint8_t* __restrict__ Vw,
                                                                const float* __restrict__ Pw, float* __restrict__ out,
                                                                int64_t ld_out, int N, int NT, float inv_sig2,
                                                                int terms, long long* __restrict__ trace) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // the 128B swizzle is anchored at 1024-byte boundaries of the shared window
    TcSmem& sm = *reinterpret_cast<TcSmem*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
    const int qt = blockIdx.x, b = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t tile0 = (size_t)b * NT;
    // optional timeline of CTA (0,0) (PDSC_ATTN_TRACE): trace[j*8 + event] = SM clock
    const bool tracing = trace != nullptr && qt == 0 && b == 0;
#define TC_TRACE(j, ev) do { if (tracing) trace[(j) * 8 + (ev)] = clock64(); } while (0)

    if (threadIdx.x == 0) {
        mbar_init(&sm.q_full, 1); mbar_init(&sm.k_full, 1); mbar_init(&sm.k_empty, 1);
        mbar_init(&sm.v_full, 1); mbar_init(&sm.v_empty, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&sm.s_full[i], 1); mbar_init(&sm.pts_full[i], 1); mbar_init(&sm.pv_done[i], 1);
            mbar_init(&sm.p_full[i][0], 4); mbar_init(&sm.p_full[i][1], 4);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(&sm.tmem_base, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = sm.tmem_base;
    const uint32_t TM_S = 0, TM_O = 256;

    if (warp == 0) {
        // ============================== producer: Q, K tiles, key points ==============================
        if (lane == 0) {
            mbar_arrive_expect_tx(&sm.q_full, TC_TILE_BYTES);
            bulk_tile(sm.q, Qw + (tile0 + qt) * TC_TILE_BYTES, &sm.q_full);
            for (int j = 0; j < NT; ++j) {
                const int buf = j & 1;
                if (j > 0) mbar_wait(&sm.k_empty, (j - 1) & 1);          // S(j-1) has consumed the K buffer
                TC_TRACE(j, 0);
                mbar_arrive_expect_tx(&sm.k_full, TC_TILE_BYTES);
                bulk_tile(sm.k, Kw + (tile0 + j) * TC_TILE_BYTES, &sm.k_full);
                if (j >= 2) {                                            // softmax(j-2) is done with pts[buf]
                    mbar_wait(&sm.p_full[buf][0], ((j - 2) >> 1) & 1);
                    mbar_wait(&sm.p_full[buf][1], ((j - 2) >> 1) & 1);
                }
                mbar_arrive_expect_tx(&sm.pts_full[buf], TC_PTS_BYTES);
                bulk_g2s(sm.pts[buf], Pw + (tile0 + j) * TC_PTS_FLOATS, TC_PTS_BYTES, &sm.pts_full[buf]);
            }
        }
    } else if (warp == 2) {
        // ============================== producer: V^T tiles (own warp: never waits behind K) ==========
        if (lane == 0) {
            for (int j = 0; j < NT; ++j) {
                if (j > 0) mbar_wait(&sm.v_empty, (j - 1) & 1);          // PV(j-1) has consumed the V buffer
                TC_TRACE(j, 1);
                mbar_arrive_expect_tx(&sm.v_full, TC_TILE_BYTES);
                bulk_tile(sm.v, Vw + (tile0 + j) * TC_TILE_BYTES, &sm.v_full);
            }
        }
    } else if (warp == 1) {
        // ============================== MMA issuer ==============================
        if (lane == 0) {
            const uint32_t q_addr = smem_u32(sm.q), k_addr = smem_u32(sm.k), v_addr = smem_u32(sm.v);
            mbar_wait(&sm.q_full, 0);
            mbar_wait(&sm.k_full, 0);
            tc_fence_after();
            issue_qk(tmem + TM_S, q_addr, k_addr, terms);
            umma_commit(&sm.s_full[0]);
            umma_commit(&sm.k_empty);
            for (int j = 0; j < NT; ++j) {
                const int buf = j & 1;
                if (j + 1 < NT) {
                    mbar_wait(&sm.k_full, (j + 1) & 1);
                    TC_TRACE(j + 1, 2);
                    tc_fence_after();
                    issue_qk(tmem + TM_S + (uint32_t)(buf ^ 1) * 128u, q_addr, k_addr, terms);
                    umma_commit(&sm.s_full[buf ^ 1]);
                    umma_commit(&sm.k_empty);
                }
                mbar_wait(&sm.v_full, j & 1);
#pragma unroll
                for (int g = 0; g < 2; ++g) {
                    mbar_wait(&sm.p_full[buf][g], (j >> 1) & 1);
                    if (g == 0) TC_TRACE(j, 3);
                    tc_fence_after();
                    issue_pv_half(tmem + TM_O + (uint32_t)g * 128u, tmem + TM_S + (uint32_t)buf * 128u + (uint32_t)g * 64u,
                                  v_addr + (uint32_t)g * TC_BLOCK_BYTES, terms, j > 0 ? 1u : 0u);
                    umma_commit(&sm.pv_done[g]);
                }
                umma_commit(&sm.v_empty);
            }
        }
    } else {
        // ============================== softmax warpgroups ==============================
        const int wg = (warp - 3) >> 2;              // keys [64 wg, 64 wg + 64) of every tile
        const int quarter = warp & 3;                // TMEM lane quarter this warp may access
        const int row = quarter * 32 + lane;         // query row inside the CTA tile
        const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
        const int qrow = min(qt * TC_BM + row, N - 1);
        const float* qp = Pw + (tile0 + (qrow >> 7)) * TC_PTS_FLOATS + ((qrow & 127) >> 1) * 12 + (qrow & 1);
        const float2 qsx = make_float2(qp[0], qp[0]), qsy = make_float2(qp[2], qp[2]), qsz = make_float2(qp[4], qp[4]);
        const float2 qtx = make_float2(qp[6], qp[6]), qty = make_float2(qp[8], qp[8]), qtz = make_float2(qp[10], qp[10]);
        const float2 neg_inv2 = make_float2(-inv_sig2, -inv_sig2), one2 = make_float2(1.0f, 1.0f);
        const uint32_t o_own = tmem + lane_base + TM_O + (uint32_t)wg * 128u;      // this warpgroup's accumulator
        float m = -INFINITY, l = 0.f;
        const bool tr = threadIdx.x == 96;

        for (int j = 0; j < NT; ++j) {
            const int buf = j & 1;
            const uint32_t par = (j >> 1) & 1;
            mbar_wait(&sm.pts_full[buf], par);
            mbar_wait(&sm.s_full[buf], par);
            if (tr) TC_TRACE(j, 4);
            tc_fence_after();
            const uint32_t s_col = tmem + lane_base + TM_S + (uint32_t)buf * 128u + (uint32_t)wg * 64u;
            const int nvalid = N - j * TC_BN;        // keys of this tile that exist
            float2 x[32];                            // 64 keys as 32 pairs
            float tmax = -INFINITY;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                uint32_t r[32];
                tmem_ld32(s_col + h * 32, r);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int kp = wg * 32 + h * 16 + i;     // key pair inside the tile
                    const float4 pa = sm.pts[buf][kp][0], pb = sm.pts[buf][kp][1], pc = sm.pts[buf][kp][2];
                    float2 dx = __fadd2_rn(qsx, make_float2(-pa.x, -pa.y));
                    float2 dy = __fadd2_rn(qsy, make_float2(-pa.z, -pa.w));
                    float2 dz = __fadd2_rn(qsz, make_float2(-pb.x, -pb.y));
                    float2 a2 = __ffma2_rn(dz, dz, __ffma2_rn(dy, dy, __fmul2_rn(dx, dx)));
                    const float2 ds = make_float2(fast_sqrt(a2.x), fast_sqrt(a2.y));
                    dx = __fadd2_rn(qtx, make_float2(-pb.z, -pb.w));
                    dy = __fadd2_rn(qty, make_float2(-pc.x, -pc.y));
                    dz = __fadd2_rn(qtz, make_float2(-pc.z, -pc.w));
                    a2 = __ffma2_rn(dz, dz, __ffma2_rn(dy, dy, __fmul2_rn(dx, dx)));
                    const float2 dt = make_float2(fast_sqrt(a2.x), fast_sqrt(a2.y));
                    const float2 c = __fadd2_rn(ds, make_float2(-dt.x, -dt.y));
                    float2 w = __ffma2_rn(__fmul2_rn(c, c), neg_inv2, one2);
                    w.x = fmaxf(w.x, 0.0f);
                    w.y = fmaxf(w.y, 0.0f);
                    const float2 xv = __fmul2_rn(w, make_float2(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1])));
                    x[h * 16 + i] = xv;
                    tmax = fmaxf(tmax, fmaxf(xv.x, xv.y));
                }
            }
            if (nvalid < TC_BN) {                    // last, partial tile: keys past N do not exist
                tmax = -INFINITY;
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const int key = wg * 64 + 2 * i;
                    if (key >= nvalid) x[i].x = -INFINITY;
                    if (key + 1 >= nvalid) x[i].y = -INFINITY;
                    tmax = fmaxf(tmax, fmaxf(x[i].x, x[i].y));
                }
            }
            if (tr) TC_TRACE(j, 5);
            // Lazy rescale: keep the old reference max unless this warpgroup's row max grew by more
            // than 2^8.  (m stays -inf while every key seen so far was past N: only warpgroup 1 of a
            // one-tile problem; then p = exp2(-inf - (-inf)) must not be evaluated.)
            if (j > 0) {
                // PV_g(j-1) has finished accumulating into O_g.  Waited for on every tile (normally
                // already complete) so that no waiter ever lags two phases behind the barrier.
                mbar_wait(&sm.pv_done[wg], (j - 1) & 1);
                if (tr) TC_TRACE(j, 6);
                const bool need = tmax > m + TC_RESCALE_LOG2;
                if (__any_sync(0xffffffffu, need)) {
                    tc_fence_after();
                    const float f = need ? fast_exp2(m - tmax) : 1.0f;   // m = -inf: f = 0, O_g was 0
                    if (need) { m = tmax; l *= f; }
#pragma unroll
                    for (int h = 0; h < 4; ++h) {
                        uint32_t r[32];
                        tmem_ld32(o_own + h * 32, r);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * f);
                        tmem_st32(o_own + h * 32, r);
                    }
                    tmem_st_wait();
                }
            } else {
                m = tmax;
            }
            // P = exp2(x - m), split into fp16 hi/lo, packed two keys per TMEM column
            const float m_use = (m == -INFINITY) ? 0.0f : m;             // all-masked rows: p = exp2(-inf) = 0
            const float2 neg_m = make_float2(-m_use, -m_use);
            float2 psum = make_float2(0.f, 0.f);
#pragma unroll
            for (int g = 0; g < 2; ++g) {
                uint32_t hi[16], lo[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float2 e = __fadd2_rn(x[g * 16 + i], neg_m);
                    const float2 p = make_float2(fast_exp2(e.x), fast_exp2(e.y));
                    psum = __fadd2_rn(psum, p);
                    hi[i] = pack_half2(p.x, p.y);
                    const float2 back = unpack_half2(hi[i]);
                    const float2 rem = __fadd2_rn(p, make_float2(-back.x, -back.y));
                    lo[i] = pack_half2(rem.x, rem.y);
                }
                tmem_st16(s_col + g * 16, hi);
                if (terms > 1) tmem_st16(s_col + 32 + g * 16, lo);
            }
            l += psum.x + psum.y;
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.p_full[buf][wg]);
            if (tr) TC_TRACE(j, 7);
        }
        // ---- epilogue: combine the two warpgroups' partial softmaxes, msg = O / L ----
        sm.mfin[wg][row] = m;
        sm.lsum[wg][row] = l;
        mbar_wait(&sm.pv_done[0], (NT - 1) & 1);
        mbar_wait(&sm.pv_done[1], (NT - 1) & 1);
        tc_fence_after();
        named_bar_sync(1, 256);
        const float m0 = sm.mfin[0][row], m1 = sm.mfin[1][row];
        const float M = fmaxf(m0, m1);
        const float a0 = (m0 == -INFINITY) ? 0.f : fast_exp2(m0 - M), a1 = (m1 == -INFINITY) ? 0.f : fast_exp2(m1 - M);
        const float inv_l = 1.0f / (sm.lsum[0][row] * a0 + sm.lsum[1][row] * a1);
        const float c0 = a0 * inv_l, c1 = a1 * inv_l;
        const int grow = qt * TC_BM + row;
        const uint32_t o_col = tmem + lane_base + TM_O + (uint32_t)wg * 64u;     // output channels [64 wg, +64)
        float* orow = out + ((size_t)b * N + min(grow, N - 1)) * ld_out + wg * 64;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            uint32_t r0[32], r1[32];
            tmem_ld32(o_col + h * 32, r0);
            tmem_ld32(o_col + 128 + h * 32, r1);
            tmem_ld_wait();
            if (grow < N) {
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    float4 o;
                    // c_g = 0 when warpgroup g saw no valid key: its accumulator may hold anything
                    o.x = (c0 != 0.f ? __uint_as_float(r0[i]) * c0 : 0.f) + (c1 != 0.f ? __uint_as_float(r1[i]) * c1 : 0.f);
                    o.y = (c0 != 0.f ? __uint_as_float(r0[i + 1]) * c0 : 0.f) + (c1 != 0.f ? __uint_as_float(r1[i + 1]) * c1 : 0.f);
                    o.z = (c0 != 0.f ? __uint_as_float(r0[i + 2]) * c0 : 0.f) + (c1 != 0.f ? __uint_as_float(r1[i + 2]) * c1 : 0.f);
                    o.w = (c0 != 0.f ? __uint_as_float(r0[i + 3]) * c0 : 0.f) + (c1 != 0.f ? __uint_as_float(r1[i + 3]) * c1 : 0.f);
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
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(attn_prep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 129 * 4);
        cudaFuncSetAttribute(attn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TcSmem) + 1024);
        attr_set = true;
    }
    const float qscale = 1.4426950408889634f / sqrtf((float)C);
    attn_prep_kernel<<<dim3(NT, B), 256, 128 * 129 * 4, st>>>(q, k, v, ld, src, tgt, N, NT, qscale, Qw, Kw, Vw, Pw);
    const float inv_sig2 = 1.0f / (sigma_spat * sigma_spat);
    const int terms = mode == PDSC_ATTN_SPLIT16 ? 3 : 1;
    long long* trace = nullptr;
    const char* trace_path = getenv("PDSC_ATTN_TRACE");     // debugging aid: timeline of CTA (0,0)
    if (trace_path) { cudaMalloc(&trace, sizeof(long long) * 8 * NT); cudaMemset(trace, 0, sizeof(long long) * 8 * NT); }
    attn_tc_kernel<<<dim3(NT, B), TC_THREADS, sizeof(TcSmem) + 1024, st>>>(Qw, Kw, Vw, Pw, out, ld_out, N, NT, inv_sig2, terms, trace);
    if (trace) {
        cudaStreamSynchronize(st);
        long long* h = (long long*)malloc(sizeof(long long) * 8 * NT);
        cudaMemcpy(h, trace, sizeof(long long) * 8 * NT, cudaMemcpyDeviceToHost);
        FILE* f = fopen(trace_path, "w");
        if (f) {
            for (int j = 0; j < NT; ++j) {
                for (int e = 0; e < 8; ++e) fprintf(f, "%lld%c", h[j * 8 + e] ? h[j * 8 + e] - h[0] : -1, e == 7 ? '\n' : ' ');
            }
            fclose(f);
        }
        free(h);
        cudaFree(trace);
    }
    count_launches(2);
    return check_launch("nonlocal_attn(tcgen05)");
}

}  // namespace pdsc
