// Fused channel-last pointwise MLP chains on the tensor cores (tcgen05, fp16 hi/lo split, fp32 accumulate).
// The 1x1-conv stacks of the reference -- PointCN (Conv1d+BN+ReLU, models/PointDSC.py:54-62) and
// NonLocalBlock.fc_message (Conv+BN+ReLU, Conv+BN+ReLU, Conv, :12-18) with the residual add (:44) -- are
// chains of tiny GEMMs (K, N <= 128) over B*N rows: as separate library GEMMs every activation makes an HBM
// round trip (plus one more for the operand packing).  Here one CTA keeps a 128-row tile on chip through the
// whole chain:
//     x (+add_in) -> [ W_0, b_0, ReLU? ] -> ... -> [ W_{L-1}, b_{L-1}, ReLU? ] (+add_out) -> y
//   threads t and t+128 <-> row t (alternate 32-column chunks, so that every scheduler has two warps to pick from):
//   they load the fp32 row, split it into fp16 hi + lo and store both to TMEM as the A operand;
//   an elected lane issues hi*hi + hi*lo + lo*hi tcgen05.mma against the pre-packed weights in shared memory
//   (K-major, 128B swizzle, BatchNorm folded); the fp32 accumulator comes back through tcgen05.ld for bias / ReLU
//   and is split again as the next layer's A operand.  Nothing but x and y touches HBM.
// Accuracy is that of the three-term split (22 mantissa bits per operand), same as pdsc_split16_rows + fp16 GEMM.
#include "common.cuh"
#include "sm100.cuh"

namespace pdsc {

using namespace sm100;

constexpr int ML_MAX_LAYERS = 4;
#ifndef ML_DIRECT_GROUPS
#define ML_DIRECT_GROUPS 2
#endif
constexpr int ML_THREADS = 256;     // two threads per row: thread t and t + 128 share row t & 127 and take alternate 32-column chunks
constexpr uint32_t ML_A_LO = 64;      // TMEM columns: A hi [0,64), A lo [64,128), D [128,256)
constexpr uint32_t ML_D = 128;

struct MlpParams {
    const float* x; const float* add_in; const float* add_out;
    float* y; uint8_t* y_packed;
    const uint8_t* wpack; const float* bias;
    long long rows;
    int n_layers;
    int K[ML_MAX_LAYERS], N[ML_MAX_LAYERS], relu[ML_MAX_LAYERS];
    uint32_t w_off[ML_MAX_LAYERS];     // byte offset of the layer's hi image inside wpack; lo image follows it
    uint32_t b_off[ML_MAX_LAYERS];     // float offset of the layer's bias
    uint32_t w_bytes, n_bias;
    int Kp;                            // row length (halves) of the packed output
    int32_t* overflow;                 // caller's operand-range flag (common.cuh), may be null
    const float* res; int res_layer;   // mlp_chain_direct_kernel: residual [rows, N[res_layer]] added to the output of that layer
};

__device__ __forceinline__ uint64_t ml_uniform64(uint64_t v) {   // broadcast from lane 0: provably warp-uniform
    const uint32_t lo = __shfl_sync(0xffffffffu, (uint32_t)v, 0), hi = __shfl_sync(0xffffffffu, (uint32_t)(v >> 32), 0);
    return ((uint64_t)hi << 32) | lo;
}
__device__ __forceinline__ void ml_split_pair(float a, float b, uint32_t& hi, uint32_t& lo, uint32_t& bad) {
    hi = pack_half2(a, b);
    bad |= half2_not_finite(hi);
    const float2 back = unpack_half2(hi);
    lo = pack_half2(a - back.x, b - back.y);
}

constexpr uint32_t ML_PITCH = 528;     // bytes between staged rows: 512 + 16, so that row-per-thread LDS.128 are conflict free
constexpr uint32_t ML_STAGE_BYTES = 128 * ML_PITCH;

// Stages rows [row0, row0 + nrows) of a row-major fp32 [rows, width] matrix into shared memory (padded pitch) with
// 16-byte cp.async copies: global reads are warp-coalesced, completion is tracked per thread in commit groups.
__device__ __forceinline__ void ml_stage_rows(uint8_t* dst, const float* src, long long row0, int nrows, int width) {
    const int sh = width == 128 ? 5 : 4;               // log2 of the 16-byte chunks per row (width is 128 or 64)
    const int total = nrows << sh;
    const float* base = src + row0 * width;
#pragma unroll 4
    for (int idx = threadIdx.x; idx < total; idx += ML_THREADS) {
        const int r = idx >> sh, c = idx & ((1 << sh) - 1);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst + (uint32_t)r * ML_PITCH + (uint32_t)c * 16u)),
                     "l"(base + (size_t)r * width + c * 4)
                     : "memory");
    }
}
__device__ __forceinline__ void ml_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void ml_wait_group() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__global__ void __launch_bounds__(ML_THREADS, 1) mlp_chain_kernel(const MlpParams p) {
    uint32_t bad = 0;                                       // operand-range test of every split (common.cuh)
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* wsm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    float* bsm = reinterpret_cast<float*>(wsm + p.w_bytes);
    uint8_t* xs = wsm + p.w_bytes + 1024;                  // staged input rows
    uint8_t* as = xs + ML_STAGE_BYTES;                     // staged residual rows (when add_out), then the fp32 output rows
    __shared__ uint64_t w_bar, mma_bar;
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        mbar_init(&w_bar, 1);
        mbar_init(&mma_bar, 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(&tmem_base_s, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;
    const uint32_t lane_addr = (uint32_t)((warp & 3) * 32) << 16;      // a warp reaches the TMEM lanes of its quadrant (warp % 4)
    const int rt = threadIdx.x & 127, half = threadIdx.x >> 7;         // row inside the tile, which chunks of it
    const long long n_tiles = (p.rows + 127) / 128;
    const int K0 = p.K[0], NL = p.N[p.n_layers - 1];
    // weights: one bulk copy per CTA (they stay resident while the CTA walks its row tiles); first tile's rows
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(&w_bar, p.w_bytes);
        for (uint32_t off = 0; off < p.w_bytes; off += 32768u) {
            const uint32_t n = min(32768u, p.w_bytes - off);
            bulk_g2s(wsm + off, p.wpack + off, n, &w_bar);
        }
    }
    // commit groups per tile, oldest first: X(tile) | A(tile) (residual, only with add_out) | X(next tile) ...
    {
        const long long row0 = (long long)blockIdx.x * 128;
        const int nrows = (int)min(128LL, p.rows - row0);
        ml_stage_rows(xs, p.x, row0, nrows, K0);
        ml_commit();
        if (p.add_out) { ml_stage_rows(as, p.add_out, row0, nrows, NL); ml_commit(); }
    }
    for (uint32_t i = threadIdx.x; i < p.n_bias; i += ML_THREADS) bsm[i] = p.bias[i];
    __syncthreads();
    mbar_wait_role(&w_bar, 0);

    uint32_t mma_phase = 0;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long long row = tile * 128 + rt;
        const bool live = row < p.rows;
        const long long rrow = live ? row : p.rows - 1;
        const long long next = tile + gridDim.x;
        // ---- input row (staged in shared memory) -> A operand (hi, lo) in TMEM ----
        if (p.add_out) ml_wait_group<1>(); else ml_wait_group<0>();
        __syncthreads();               // X(tile) has landed for every thread's copies
        {
            const float4* xr = reinterpret_cast<const float4*>(xs + (uint32_t)rt * ML_PITCH);
            const float4* ar = p.add_in ? reinterpret_cast<const float4*>(p.add_in + rrow * K0) : nullptr;
            for (int c = half; c < K0 / 32; c += 2) {
                uint32_t hi[16], lo[16];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    float4 v = xr[c * 8 + i];
                    if (ar) { const float4 a = __ldg(ar + c * 8 + i); v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w; }
                    ml_split_pair(v.x, v.y, hi[2 * i], lo[2 * i], bad);
                    ml_split_pair(v.z, v.w, hi[2 * i + 1], lo[2 * i + 1], bad);
                }
                tmem_st16(tmem + lane_addr + (uint32_t)c * 16u, hi);
                tmem_st16(tmem + lane_addr + ML_A_LO + (uint32_t)c * 16u, lo);
            }
        }
        for (int l = 0; l < p.n_layers; ++l) {
            const int K = p.K[l], N = p.N[l];
            tmem_st_wait();
            tc_fence_before();
            __syncthreads();           // every row's A operand is in TMEM, and everyone is done reading D (and xs)
            if (warp == 0) {
                tc_fence_after();
                const uint32_t idesc = umma_idesc_f16(128, N);
                const uint32_t blk = (uint32_t)N * 128u;                  // bytes of one 64-channel B block
                const uint32_t lo_off = (uint32_t)(K / 64) * blk;         // lo image follows the hi image
                const uint64_t wd = ml_uniform64(umma_desc_sw128_kmajor(smem_u32(wsm + p.w_off[l])));
                uint32_t acc = 0;
                // cross terms first (term order 1, 2, 0 = hi*lo, lo*hi, hi*hi): the TMEM accumulator truncates once per MMA at
                // the magnitude of the running sum, so the small terms go in while that sum is still small
                for (int ti = 0; ti < 3; ++ti) {
                    const int term = ti == 2 ? 0 : ti + 1;
                    const uint32_t a0 = tmem + (term == 2 ? ML_A_LO : 0u);
                    const uint32_t b0 = term == 1 ? lo_off : 0u;
                    for (int kk = 0; kk < K / 16; ++kk) {
                        const uint64_t db = wd + (uint64_t)((b0 + (uint32_t)(kk >> 2) * blk + (uint32_t)(kk & 3) * 32u) >> 4);
                        umma_ts_elect(tmem + ML_D, a0 + (uint32_t)kk * 8u, db, idesc, acc);
                        acc = 1;
                    }
                }
                umma_commit_elect(&mma_bar);
            }
            if (l == 0) {
                // the input stage is free (all rows were read before the barrier above): prefetch the next tile
                if (next < n_tiles) ml_stage_rows(xs, p.x, next * 128, (int)min(128LL, p.rows - next * 128), K0);
                ml_commit();           // (an empty group when there is no next tile keeps the group count uniform)
            }
            const bool last = l + 1 == p.n_layers;
            mbar_wait(&mma_bar, mma_phase);
            mma_phase ^= 1;
            tc_fence_after();
            if (last && p.add_out) { ml_wait_group<1>(); __syncthreads(); }   // A(tile) has landed; X(next) may still fly
            const float* bias = bsm + p.b_off[l];
            for (int c = half; c < N / 32; c += 2) {
                uint32_t r[32];
                tmem_ld32(tmem + lane_addr + ML_D + (uint32_t)c * 32u, r);
                tmem_ld_wait();
                float v[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    v[i] = __uint_as_float(r[i]) + bias[c * 32 + i];
                    if (p.relu[l]) v[i] = fmaxf(v[i], 0.f);
                }
                if (!last) {
                    uint32_t hi[16], lo[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) ml_split_pair(v[2 * i], v[2 * i + 1], hi[i], lo[i], bad);
                    tmem_st16(tmem + lane_addr + (uint32_t)c * 16u, hi);
                    tmem_st16(tmem + lane_addr + ML_A_LO + (uint32_t)c * 16u, lo);
                } else {
                    // residual add in place in the stage; the finished fp32 row leaves through per-row bulk stores
                    float4* orow = reinterpret_cast<float4*>(as + (uint32_t)rt * ML_PITCH) + c * 8;
                    if (p.add_out) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const float4 a = orow[i];
                            v[4 * i] += a.x; v[4 * i + 1] += a.y; v[4 * i + 2] += a.z; v[4 * i + 3] += a.w;
                        }
                    }
                    if (p.y) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) orow[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
                    }
                    if (live && p.y_packed) {
                        // row layout of pdsc_split16_rows: [hi(N) | hi(N) | lo(N) | 1 1 0 ..]
                        uint8_t* pr = p.y_packed + (size_t)row * p.Kp * 2;
                        uint32_t hi[16], lo[16];
#pragma unroll
                        for (int i = 0; i < 16; ++i) ml_split_pair(v[2 * i], v[2 * i + 1], hi[i], lo[i], bad);
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const uint4 h4 = make_uint4(hi[4 * i], hi[4 * i + 1], hi[4 * i + 2], hi[4 * i + 3]);
                            const uint4 l4 = make_uint4(lo[4 * i], lo[4 * i + 1], lo[4 * i + 2], lo[4 * i + 3]);
                            *reinterpret_cast<uint4*>(pr + (c * 32 + i * 8) * 2) = h4;
                            *reinterpret_cast<uint4*>(pr + (N + c * 32 + i * 8) * 2) = h4;
                            *reinterpret_cast<uint4*>(pr + (2 * N + c * 32 + i * 8) * 2) = l4;
                        }
                        if (c == 0) {
                            // tail: 1 1 0 0 0 ... (Kp - 3N halves, a multiple of 8)
                            for (int h8 = 0; h8 < (p.Kp - 3 * N) / 8; ++h8)
                                *reinterpret_cast<uint4*>(pr + (3 * N + h8 * 8) * 2) = make_uint4(h8 == 0 ? 0x3c003c00u : 0u, 0u, 0u, 0u);
                        }
                    }
                }
            }
        }
        if (p.y || p.add_out) {
            __syncthreads();                   // every row of the stage is final (and the residual has been consumed)
            if (p.y) {
                // stage -> global, warp-coalesced: one instruction writes one whole row
                const int nrows = (int)min(128LL, p.rows - tile * 128);
                const int sh = NL == 128 ? 5 : 4, total = nrows << sh;
                float4* ybase = reinterpret_cast<float4*>(p.y + tile * 128 * NL);
#pragma unroll 4
                for (int idx = threadIdx.x; idx < total; idx += ML_THREADS) {
                    const int r = idx >> sh, c = idx & ((1 << sh) - 1);
                    ybase[idx] = *reinterpret_cast<const float4*>(as + (uint32_t)r * ML_PITCH + (uint32_t)c * 16u);
                }
                __syncthreads();               // the stage has been read
            }
            if (p.add_out) {
                if (next < n_tiles) ml_stage_rows(as, p.add_out, next * 128, (int)min(128LL, p.rows - next * 128), NL);
                ml_commit();
            }
        }
    }
    ml_wait_group<0>();
    if (bad && p.overflow) atomicOr(p.overflow, 1);
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 256);
}

// ------------------------------------------------------------------------------------------------
// The same chain with the rows moved straight between global memory and registers, shared memory = weights only, so that a
// chain of four layers fits: fc_message + residual of one encoder layer AND the PointCN of the next in one launch
// (models/PointDSC.py:12-18,44 then :54-62) -- the sum  feat + fc_message(msg)  is consumed on chip and never written.
//     x -> L_0 .. L_{r} (+ res) -> L_{r+1} .. L_{n-1} -> y
// A thread still owns one 128-byte piece (32 channels) of a row at a time, which as a plain load would touch 32 lines per
// warp instruction (round 1 measured that: slower than staging).  Here the eight lanes of a group read ONE whole line per
// instruction -- lane k the k-th float4 of the row of lane 8g + j in instruction j -- and an 8 x 8 register transpose
// (warp_transpose8) hands every lane its own piece; stores go the other way round.
__device__ __forceinline__ void ml_load_piece(const float* src, int width, long long row0, long long rows, int c, int lane, uint32_t (&r)[32]) {
    const int g8 = lane & ~7, k = lane & 7;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const long long row = min(row0 + g8 + j, rows - 1);
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(src + row * width + c * 32) + k);
        r[4 * j] = v.x; r[4 * j + 1] = v.y; r[4 * j + 2] = v.z; r[4 * j + 3] = v.w;
    }
}
__device__ __forceinline__ void ml_store_piece(float* dst, int width, long long row0, long long rows, int c, int lane, uint32_t (&r)[32]) {
    const int g8 = lane & ~7, k = lane & 7;
    warp_transpose8(r, k);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const long long row = row0 + g8 + j;
        if (row < rows)
            *(reinterpret_cast<uint4*>(dst + row * width + c * 32) + k) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
    }
}

// GROUPS = 1: 8 warps, the next tile's input pieces and this tile's residual pieces are fetched into registers ahead of use
// (255 registers).  GROUPS = 2: two independent 8-warp groups per CTA, each walking its own tiles with its own 256 TMEM
// columns and the weights shared in shared memory; no register prefetch (128 registers) -- while one group waits for a
// load, an MMA or a barrier the other one runs.  (prefetch.global.L2 of the next tile's input and this tile's residual at the
// top of a tile was measured: 9.976 vs 9.963 ms per forward at 16 x N=5000, 3.632 vs 3.602 at 64 x N=1000 -- not kept.)
template <int GROUPS>
__global__ void __launch_bounds__(ML_THREADS * GROUPS, 1) mlp_chain_direct_kernel(const MlpParams p) {
    constexpr bool PREFETCH = GROUPS == 1;
    uint32_t bad = 0;                                       // operand-range test of every split (common.cuh)
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* wsm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    float* bsm = reinterpret_cast<float*>(wsm + p.w_bytes);
    __shared__ uint64_t w_bar, mma_bar[GROUPS];
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int gi = threadIdx.x / ML_THREADS, gt = threadIdx.x % ML_THREADS;      // group, thread inside the group
    auto group_sync = [&]() { if (GROUPS == 1) __syncthreads(); else named_bar_sync(1 + gi, ML_THREADS); };

    if (threadIdx.x == 0) {
        mbar_init(&w_bar, 1);
        for (int i = 0; i < GROUPS; ++i) mbar_init(&mma_bar[i], 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(&tmem_base_s, 256 * GROUPS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s + 256u * gi;
    const uint32_t lane_addr = (uint32_t)((warp & 3) * 32) << 16;      // a warp reaches the TMEM lanes of its quadrant (warp % 4)
    const int half = gt >> 7;                                          // which 32-channel pieces of its row this thread takes
    const int wrow0 = (warp & 3) * 32;                                 // first row of the tile this warp's lanes own
    const int k8 = lane & 7;
    const long long n_tiles = (p.rows + 127) / 128;
    const long long tile0 = (long long)blockIdx.x * GROUPS + gi, tile_step = (long long)gridDim.x * GROUPS;
    const int K0 = p.K[0], NL = p.N[p.n_layers - 1], NR = p.N[p.res_layer];
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(&w_bar, p.w_bytes);
        for (uint32_t off = 0; off < p.w_bytes; off += 32768u) {
            const uint32_t n = min(32768u, p.w_bytes - off);
            bulk_g2s(wsm + off, p.wpack + off, n, &w_bar);
        }
    }
    // pieces c = half, half + 2 of the first tile's rows (in flight while the weights arrive)
    uint32_t xin[2][32], rin[2][32];
    if (PREFETCH && tile0 < n_tiles) {
#pragma unroll
        for (int ci = 0; ci < 2; ++ci)
            if (half + 2 * ci < K0 / 32) ml_load_piece(p.x, K0, tile0 * 128 + wrow0, p.rows, half + 2 * ci, lane, xin[ci]);
    }
    for (uint32_t i = threadIdx.x; i < p.n_bias; i += ML_THREADS * GROUPS) bsm[i] = p.bias[i];
    __syncthreads();
    mbar_wait_role(&w_bar, 0);

    uint32_t mma_phase = 0;
    for (long long tile = tile0; tile < n_tiles; tile += tile_step) {
        const long long row0 = tile * 128 + wrow0;
        const long long next = tile + tile_step;
        // ---- input pieces -> A operand (hi, lo) in TMEM ----
#pragma unroll
        for (int ci = 0; ci < 2; ++ci) {
            const int c = half + 2 * ci;
            if (c < K0 / 32) {
                if (!PREFETCH) ml_load_piece(p.x, K0, row0, p.rows, c, lane, xin[ci]);
                warp_transpose8(xin[ci], k8);
                uint32_t hi[16], lo[16];
#pragma unroll
                for (int i = 0; i < 16; ++i)
                    ml_split_pair(__uint_as_float(xin[ci][2 * i]), __uint_as_float(xin[ci][2 * i + 1]), hi[i], lo[i], bad);
                tmem_st16(tmem + lane_addr + (uint32_t)c * 16u, hi);
                tmem_st16(tmem + lane_addr + ML_A_LO + (uint32_t)c * 16u, lo);
            }
        }
        if (PREFETCH) {
            // the residual pieces of this tile and the input pieces of the next one fly while the chain runs
#pragma unroll
            for (int ci = 0; ci < 2; ++ci) {
                if (p.res && half + 2 * ci < NR / 32) ml_load_piece(p.res, NR, row0, p.rows, half + 2 * ci, lane, rin[ci]);
                if (next < n_tiles && half + 2 * ci < K0 / 32) ml_load_piece(p.x, K0, next * 128 + wrow0, p.rows, half + 2 * ci, lane, xin[ci]);
            }
        }
        for (int l = 0; l < p.n_layers; ++l) {
            const int K = p.K[l], N = p.N[l];
            tmem_st_wait();
            tc_fence_before();
            group_sync();              // every row's A operand is in TMEM, and everyone is done reading D
            if ((warp & 7) == 0) {
                tc_fence_after();
                const uint32_t idesc = umma_idesc_f16(128, N);
                const uint32_t blk = (uint32_t)N * 128u;                  // bytes of one 64-channel B block
                const uint32_t lo_off = (uint32_t)(K / 64) * blk;         // lo image follows the hi image
                const uint64_t wd = ml_uniform64(umma_desc_sw128_kmajor(smem_u32(wsm + p.w_off[l])));
                uint32_t acc = 0;
                for (int ti = 0; ti < 3; ++ti) {                          // cross terms first (see mlp_chain_kernel)
                    const int term = ti == 2 ? 0 : ti + 1;
                    const uint32_t a0 = tmem + (term == 2 ? ML_A_LO : 0u);
                    const uint32_t b0 = term == 1 ? lo_off : 0u;
                    for (int kk = 0; kk < K / 16; ++kk) {
                        const uint64_t db = wd + (uint64_t)((b0 + (uint32_t)(kk >> 2) * blk + (uint32_t)(kk & 3) * 32u) >> 4);
                        umma_ts_elect(tmem + ML_D, a0 + (uint32_t)kk * 8u, db, idesc, acc);
                        acc = 1;
                    }
                }
                umma_commit_elect(&mma_bar[gi]);
            }
            const bool last = l + 1 == p.n_layers;
            const bool add_res = p.res && l == p.res_layer;
            if (!PREFETCH && add_res) {
#pragma unroll
                for (int ci = 0; ci < 2; ++ci)
                    if (half + 2 * ci < N / 32) ml_load_piece(p.res, NR, row0, p.rows, half + 2 * ci, lane, rin[ci]);     // (lands while the MMAs run)
            }
            mbar_wait(&mma_bar[gi], mma_phase);
            mma_phase ^= 1;
            tc_fence_after();
            const float* bias = bsm + p.b_off[l];
#pragma unroll
            for (int ci = 0; ci < 2; ++ci) {
                const int c = half + 2 * ci;
                if (c >= N / 32) continue;
                uint32_t r[32];
                tmem_ld32(tmem + lane_addr + ML_D + (uint32_t)c * 32u, r);
                tmem_ld_wait();
                float v[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    v[i] = __uint_as_float(r[i]) + bias[c * 32 + i];
                    if (p.relu[l]) v[i] = fmaxf(v[i], 0.f);
                }
                if (add_res) {
                    warp_transpose8(rin[ci], k8);
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] += __uint_as_float(rin[ci][i]);
                }
                if (!last) {
                    uint32_t hi[16], lo[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) ml_split_pair(v[2 * i], v[2 * i + 1], hi[i], lo[i], bad);
                    tmem_st16(tmem + lane_addr + (uint32_t)c * 16u, hi);
                    tmem_st16(tmem + lane_addr + ML_A_LO + (uint32_t)c * 16u, lo);
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(v[i]);
                    ml_store_piece(p.y, NL, row0, p.rows, c, lane, r);
                }
            }
        }
    }
    if (bad && p.overflow) atomicOr(p.overflow, 1);
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base_s, 256 * GROUPS);
}

int launch_mlp_chain(const float* x, const float* add_in, long long rows, int n_layers, const int32_t* dims,
                     const int32_t* relu, const void* wpack, const float* bias, const float* add_out, float* y,
                     void* y_packed, int Kp, cudaStream_t st) {
    if (n_layers < 1 || n_layers > ML_MAX_LAYERS) { set_error("mlp_chain: 1..%d layers supported", ML_MAX_LAYERS); return PDSC_ERR_ARG; }
    MlpParams p;
    p.x = x; p.add_in = add_in; p.add_out = add_out; p.y = y; p.y_packed = reinterpret_cast<uint8_t*>(y_packed);
    p.wpack = reinterpret_cast<const uint8_t*>(wpack); p.bias = bias; p.rows = rows; p.n_layers = n_layers; p.Kp = Kp;
    p.overflow = overflow_flag();
    p.res = nullptr; p.res_layer = 0;
    uint32_t woff = 0, boff = 0;
    for (int l = 0; l < n_layers; ++l) {
        const int K = dims[l], N = dims[l + 1];
        if ((K != 64 && K != 128) || (N != 64 && N != 128)) { set_error("mlp_chain: layer dims must be 64 or 128 (got %d -> %d)", K, N); return PDSC_ERR_ARG; }
        p.K[l] = K; p.N[l] = N; p.relu[l] = relu[l];
        p.w_off[l] = woff; p.b_off[l] = boff;
        woff += 2u * (uint32_t)(K / 64) * (uint32_t)N * 128u;     // hi + lo images
        boff += (uint32_t)N;
    }
    p.w_bytes = woff; p.n_bias = boff;
    if (y_packed && (Kp % 8 != 0 || Kp < 3 * dims[n_layers] + 8)) { set_error("mlp_chain: bad packed row length %d", Kp); return PDSC_ERR_ARG; }
    // weights | biases (1 KB slot) | staged input rows | staged residual rows
    if (p.n_bias * sizeof(float) > 1024) { set_error("mlp_chain: too many bias entries"); return PDSC_ERR_ARG; }
    const int smem = (int)(p.w_bytes + 1024 + ML_STAGE_BYTES + ((add_out || y) ? ML_STAGE_BYTES : 0) + 1024);
    if (smem > 227 * 1024) { set_error("mlp_chain: chain does not fit in shared memory (%d bytes)", smem); return PDSC_ERR_ARG; }
    opt_in_smem<mlp_chain_kernel>(smem);
    const long long n_tiles = (rows + 127) / 128;
    long long grid = num_sms();
    if (grid > n_tiles) grid = n_tiles;
    mlp_chain_kernel<<<(int)grid, ML_THREADS, smem, st>>>(p);
    count_launches(1);
    return check_launch("mlp_chain");
}

int launch_mlp_chain_direct(const float* x, long long rows, int n_layers, const int32_t* dims, const int32_t* relu, const void* wpack,
                            const float* bias, const float* res, int res_layer, float* y, cudaStream_t st) {
    if (n_layers < 1 || n_layers > ML_MAX_LAYERS) { set_error("mlp_chain_direct: 1..%d layers supported", ML_MAX_LAYERS); return PDSC_ERR_ARG; }
    if (res && (res_layer < 0 || res_layer >= n_layers)) { set_error("mlp_chain_direct: residual layer %d outside the chain", res_layer); return PDSC_ERR_ARG; }
    MlpParams p;
    p.x = x; p.add_in = nullptr; p.add_out = nullptr; p.y = y; p.y_packed = nullptr; p.Kp = 0;
    p.wpack = reinterpret_cast<const uint8_t*>(wpack); p.bias = bias; p.rows = rows; p.n_layers = n_layers;
    p.res = res; p.res_layer = res ? res_layer : 0;
    p.overflow = overflow_flag();
    uint32_t woff = 0, boff = 0;
    for (int l = 0; l < n_layers; ++l) {
        const int K = dims[l], N = dims[l + 1];
        if ((K != 64 && K != 128) || (N != 64 && N != 128)) { set_error("mlp_chain_direct: layer dims must be 64 or 128 (got %d -> %d)", K, N); return PDSC_ERR_ARG; }
        p.K[l] = K; p.N[l] = N; p.relu[l] = relu[l];
        p.w_off[l] = woff; p.b_off[l] = boff;
        woff += 2u * (uint32_t)(K / 64) * (uint32_t)N * 128u;
        boff += (uint32_t)N;
    }
    p.w_bytes = woff; p.n_bias = boff;
    const int smem = (int)(p.w_bytes + p.n_bias * sizeof(float) + 1024);       // weights | biases
    if (smem > 227 * 1024) { set_error("mlp_chain_direct: weights do not fit in shared memory (%d bytes)", smem); return PDSC_ERR_ARG; }
    const long long n_tiles = (rows + 127) / 128;
    long long grid = num_sms();
    const int groups = (ML_DIRECT_GROUPS == 2 && n_tiles >= 2 * grid) ? 2 : 1;      // two tiles in flight per SM when there are enough
    if (groups == 2) {
        opt_in_smem<mlp_chain_direct_kernel<2>>(smem);
        mlp_chain_direct_kernel<2><<<(int)grid, ML_THREADS * 2, smem, st>>>(p);
    } else {
        if (grid > n_tiles) grid = n_tiles;
        opt_in_smem<mlp_chain_direct_kernel<1>>(smem);
        mlp_chain_direct_kernel<1><<<(int)grid, ML_THREADS, smem, st>>>(p);
    }
    count_launches(1);
    return check_launch("mlp_chain_direct");
}

}  // namespace pdsc
