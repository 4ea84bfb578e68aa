// Fused q/k/v projection + attention operand packing.
// NonLocalBlock.forward (models/PointDSC.py:29-37) projects the layer input three times (projection_q/k/v, 1x1 convs);
// the attention kernel (attn_pp.cu) wants the results as fp16 hi/lo tiles in its shared-memory image (Q and K
// row-major swizzled, V transposed).  This kernel goes from the fp32 layer input straight to those tiles:
//   * the input tile [128 rows x 128 ch] is split into fp16 hi + lo and written to shared memory as a swizzled
//     K-major tcgen05 operand;
//   * Q, K:  D[row][ch]  = X W^T  (A = input tile, B = weights);   V:  D[ch][key] = W X^T  (A = weights, B = input
//     tile), so V comes out of the tensor core already transposed -- no transpose pass;
//   * hi*hi + hi*lo + lo*hi with fp32 accumulation in TMEM (fp32-grade accuracy);
//   * the epilogue adds the bias (Q: also the log2(e)/sqrt(C) scale), splits again and assembles the finished 64 KB
//     tile image in shared memory, which leaves as ONE bulk (TMA) store.
// One CTA = one projection (blockIdx.y), walking the 128-row tiles of all pairs with two independent 8-warp groups
// (own operand/image buffer, own accumulator columns, own barriers), so that the loads of one tile overlap the
// MMAs / epilogue / store of another; the K CTAs also write the sigma-normalised key points.  Replaces: operand packing + library GEMM + attn_pp_prep_kernel.
#include "common.cuh"
#include "sm100.cuh"

#include <math.h>

namespace pdsc {

using namespace sm100;

void attn_pp_tile_pointers(void* ws, int B, int N, uint8_t** Qw, uint8_t** Kw, uint8_t** Vw, float** Pw, int** counters,
                           int* n_counters);   // attn_pp.cu
int attn_pp_tiles128(int N);

constexpr int QK_GROUP = 256;         // threads per group: two per row (thread t and t + 128 take alternate 32-column chunks of row t)
constexpr int QK_THREADS = 2 * QK_GROUP;  // two independent 8-warp groups per CTA (even / odd tiles) sharing the weights
constexpr uint32_t QK_BLK = 16384;        // one operand block: 128 rows x 64 fp16, 128-byte rows, 128B swizzle
constexpr uint32_t QK_IMG = 4 * QK_BLK;   // hi blk0 | hi blk1 | lo blk0 | lo blk1
constexpr uint32_t QK_PTS_FLOATS = 32 * 12 + 4;   // = PP_PTS_FLOATS (attn_pp.cu)

struct QkvParams {
    const float* y; const uint8_t* wpack; const float* bias; const float* src; const float* tgt;
    uint8_t* Qw; uint8_t* Kw; uint8_t* Vw; float* Pw; int* counters;
    int n_counters;
    int B, N, NT128;
    float qscale, inv_sigma;
    int32_t* overflow;                 // caller's operand-range flag (common.cuh), may be null
};

__device__ __forceinline__ uint32_t qk_swz(uint32_t r, uint32_t c) { return r * 128u + ((c ^ (r & 7u)) << 4); }
__device__ __forceinline__ uint64_t qk_uniform64(uint64_t v) {
    const uint32_t lo = __shfl_sync(0xffffffffu, (uint32_t)v, 0), hi = __shfl_sync(0xffffffffu, (uint32_t)(v >> 32), 0);
    return ((uint64_t)hi << 32) | lo;
}
__device__ __forceinline__ void qk_split_pair(float a, float b, uint32_t& hi, uint32_t& lo, uint32_t& bad) {
    hi = pack_half2(a, b);
    bad |= half2_not_finite(hi);
    const float2 back = unpack_half2(hi);
    lo = pack_half2(a - back.x, b - back.y);
}

__global__ void __launch_bounds__(QK_THREADS, 1) qkv_pack_kernel(const QkvParams p) {
    uint32_t bad = 0;                                       // operand-range test of every split (common.cuh)
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* wsm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    __shared__ float bsm[128];
    __shared__ uint64_t w_bar, mma_bars[2];
    __shared__ uint32_t tmem_base_s;
    const int grp = threadIdx.x >> 8;     // warp group: tiles grp, grp + 2, ... of this CTA's share
    const int warp = (threadIdx.x >> 5) & 7, lane = threadIdx.x & 31, gt = threadIdx.x & 255;
    const int tid = gt & 127, half = gt >> 7;       // row of the tile this thread works on in the epilogue, and which chunks of it
    const int proj = blockIdx.y;          // 0 q, 1 k, 2 v
    // per group: the input tile as a tcgen05 operand (hi/lo, swizzled); after the MMAs the same buffer receives the
    // finished output tile image
    uint8_t* yimg = wsm + QK_IMG + (uint32_t)grp * QK_IMG;
    uint8_t* img = yimg;
    uint64_t& mma_bar = mma_bars[grp];

    if (threadIdx.x == 0) {
        mbar_init(&w_bar, 1);
        mbar_init(&mma_bars[0], 1);
        mbar_init(&mma_bars[1], 1);
        fence_barrier_init();
    }
    if (threadIdx.x < 32) tmem_alloc(&tmem_base_s, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s + (uint32_t)grp * 128u;
    const uint32_t lane_addr = (uint32_t)((warp & 3) * 32) << 16;      // TMEM lanes of the warp's quadrant (warp % 4)
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(&w_bar, QK_IMG);
        bulk_g2s(wsm, p.wpack + (size_t)proj * QK_IMG, QK_IMG / 2, &w_bar);
        bulk_g2s(wsm + QK_IMG / 2, p.wpack + (size_t)proj * QK_IMG + QK_IMG / 2, QK_IMG / 2, &w_bar);
    }
    if (threadIdx.x < 128) bsm[threadIdx.x] = p.bias[proj * 128 + threadIdx.x];
    __syncthreads();
    mbar_wait_role(&w_bar, 0);

    const int N = p.N;
    const long long n_tiles = (long long)p.B * p.NT128;
    uint32_t mma_phase = 0;
    if (blockIdx.x == 0 && blockIdx.y == 0)     // arrival counters of the attention kernel start at zero (no memset on the stream)
        for (int i = threadIdx.x; i < p.n_counters; i += QK_THREADS) p.counters[i] = 0;
    for (long long tile = blockIdx.x + (long long)grp * gridDim.x; tile < n_tiles; tile += 2LL * gridDim.x) {
        const int b = (int)(tile / p.NT128), t = (int)(tile - (long long)b * p.NT128);
        const int n0 = t * 128;
        // ---- 1. input tile -> fp16 hi/lo operand image (warp-coalesced global reads, 8 rows in flight per warp) ----
        {
            const float* ybase = p.y + ((size_t)b * N + n0) * 128 + 4 * lane;
            const uint32_t blk = (uint32_t)(lane >> 4) * QK_BLK, chunk = (uint32_t)(lane & 15) >> 1, sub = (uint32_t)(lane & 1) * 8u;
            float4 v[16];                  // the whole share of this warp is requested before any of it is used
#pragma unroll
            for (int u = 0; u < 16; ++u) {
                const int r = warp + 8 * u;
                v[u] = (n0 + r < N) ? __ldg(reinterpret_cast<const float4*>(ybase + (size_t)r * 128)) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
            if (gt == 0) bulk_wait_read();     // the image this buffer held (two tiles ago) has left shared memory
            named_bar_sync(1 + grp, QK_GROUP);
#pragma unroll
            for (int u = 0; u < 16; ++u) {
                const uint32_t r = (uint32_t)(warp + 8 * u);
                uint32_t h0, l0, h1, l1;
                qk_split_pair(v[u].x, v[u].y, h0, l0, bad);
                qk_split_pair(v[u].z, v[u].w, h1, l1, bad);
                const uint32_t off = blk + qk_swz(r, chunk) + sub;
                *reinterpret_cast<uint2*>(yimg + off) = make_uint2(h0, h1);
                *reinterpret_cast<uint2*>(yimg + 2 * QK_BLK + off) = make_uint2(l0, l1);
            }
        }
        fence_proxy_async();               // st.shared operand -> tensor-core (async proxy) reads
        tc_fence_before();
        named_bar_sync(1 + grp, QK_GROUP); // operand complete; the group is done reading D of its previous tile
        // ---- 2. D = X W^T (q, k) or W X^T (v): hi*hi + hi*lo + lo*hi ----
        if (warp == 0) {
            tc_fence_after();
            const uint64_t wd = qk_uniform64(umma_desc_sw128_kmajor(smem_u32(wsm)));
            const uint64_t yd = qk_uniform64(umma_desc_sw128_kmajor(smem_u32(yimg)));
            const uint64_t da = proj < 2 ? yd : wd, db = proj < 2 ? wd : yd;
            constexpr uint32_t idesc = umma_idesc_f16(128, 128);
            uint32_t acc = 0;
#pragma unroll
            // cross terms first (hi*lo, lo*hi, then hi*hi): see mlp.cu
            for (int ti = 0; ti < 3; ++ti) {
                const int term = ti == 2 ? 0 : ti + 1;
                const uint32_t a_lo = term == 2 ? (2u * QK_BLK) >> 4 : 0u, b_lo = term == 1 ? (2u * QK_BLK) >> 4 : 0u;
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    const uint32_t koff = ((uint32_t)(kk >> 2) * QK_BLK + (uint32_t)(kk & 3) * 32u) >> 4;
                    umma_ss_elect(tmem, da + a_lo + koff, db + b_lo + koff, idesc, acc);
                    acc = 1;
                }
            }
            umma_commit_elect(&mma_bar);
        }
        if (proj == 1 && half == 0) {
            // key points of this tile for the compatibility term (sigma-normalised, packed per key pair)
            const int n = min(n0 + tid, N - 1);
            const float* s = p.src + ((size_t)b * N + n) * 3;
            const float* g = p.tgt + ((size_t)b * N + n) * 3;
            const int pr = (tid & 63) >> 1;
            float* d = p.Pw + ((size_t)tile * 2 + (tid >> 6)) * QK_PTS_FLOATS + pr * 12 + (pr >= 16 ? 4 : 0) + (tid & 1);
            d[0] = s[0] * p.inv_sigma; d[2] = s[1] * p.inv_sigma; d[4] = s[2] * p.inv_sigma;
            d[6] = g[0] * p.inv_sigma; d[8] = g[1] * p.inv_sigma; d[10] = g[2] * p.inv_sigma;
        }
        mbar_wait(&mma_bar, mma_phase);
        mma_phase ^= 1;
        tc_fence_after();
        named_bar_sync(1 + grp, QK_GROUP); // all MMAs have read the operand buffer: it may now receive the image
        // ---- 3. epilogue: bias (+ scale), split, assemble the tile image ----
        {
            const float bias_t = bsm[tid];
            const bool row_live = n0 + tid < N;          // q, k: this thread's row exists
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
                const int c = 2 * cc + half;
                uint32_t r[32];
                tmem_ld32(tmem + lane_addr + (uint32_t)c * 32u, r);
                tmem_ld_wait();
                float v[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    float x = __uint_as_float(r[i]);
                    if (proj == 2) {
                        x = (n0 + c * 32 + i < N) ? x + bias_t : 0.f;            // column = key
                    } else {
                        x = row_live ? x + bsm[c * 32 + i] : 0.f;                // column = channel
                        if (proj == 0) x *= p.qscale;
                    }
                    v[i] = x;
                }
                // destination of this thread's 4 chunks (8 halves each) inside the image
                uint32_t base, lo_off;
                if (proj == 0) { base = (uint32_t)(c >> 1) * QK_BLK + (uint32_t)tid * 128u; lo_off = 2u * QK_BLK; }
                else if (proj == 1) { base = (uint32_t)(tid >> 6) * 32768u + (uint32_t)(c >> 1) * 8192u + (uint32_t)(tid & 63) * 128u; lo_off = 16384u; }
                else { base = (uint32_t)(c >> 1) * 32768u + (uint32_t)tid * 128u; lo_off = 16384u; }
                const uint32_t r7 = (uint32_t)tid & 7u;     // (tid & 63) & 7 == tid & 7
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    uint32_t h[4], l[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) qk_split_pair(v[8 * j + 2 * e], v[8 * j + 2 * e + 1], h[e], l[e], bad);
                    const uint32_t off = base + ((((uint32_t)(c & 1) * 4u + (uint32_t)j) ^ r7) << 4);
                    *reinterpret_cast<uint4*>(img + off) = make_uint4(h[0], h[1], h[2], h[3]);
                    *reinterpret_cast<uint4*>(img + lo_off + off) = make_uint4(l[0], l[1], l[2], l[3]);
                }
            }
        }
        fence_proxy_async();               // image written with st.shared, read by the bulk store
        tc_fence_before();
        named_bar_sync(1 + grp, QK_GROUP);
        if (gt == 0) {
            uint8_t* dst = (proj == 0 ? p.Qw : proj == 1 ? p.Kw : p.Vw) + (size_t)tile * QK_IMG;
            bulk_s2g(dst, img, QK_IMG / 2);
            bulk_s2g(dst + QK_IMG / 2, img + QK_IMG / 2, QK_IMG / 2);
            bulk_commit();
        }
    }
    if (gt == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    if (bad && p.overflow) atomicOr(p.overflow, 1);
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(tmem_base_s, 256);
}

int launch_qkv_pack(const float* y, const void* wpack, const float* bias, const float* src, const float* tgt, int B, int N,
                    int C, float sigma_spat, void* ws, size_t ws_bytes, cudaStream_t st) {
    if (C != 128) { set_error("qkv_pack: C=%d unsupported (built for C=128)", C); return PDSC_ERR_ARG; }
    if (!ws || ws_bytes < attn_pp_workspace(B, N)) {
        set_error("qkv_pack: workspace too small (%zu < %zu)", ws_bytes, attn_pp_workspace(B, N));
        return PDSC_ERR_WORKSPACE;
    }
    QkvParams p;
    p.y = y; p.wpack = reinterpret_cast<const uint8_t*>(wpack); p.bias = bias; p.src = src; p.tgt = tgt;
    attn_pp_tile_pointers(ws, B, N, &p.Qw, &p.Kw, &p.Vw, &p.Pw, &p.counters, &p.n_counters);
    p.B = B; p.N = N; p.NT128 = attn_pp_tiles128(N);
    p.overflow = overflow_flag();
    p.qscale = 1.4426950408889634f / sqrtf((float)C);
    p.inv_sigma = 1.0f / sigma_spat;
    const int smem = (int)(3 * QK_IMG + 1024);   // weights + one operand/image buffer per warp group
    opt_in_smem<qkv_pack_kernel>(smem);
    const long long n_tiles = (long long)B * p.NT128;
    long long gx = num_sms() / 3;
    if (gx > n_tiles) gx = n_tiles;
    if (gx < 1) gx = 1;
    qkv_pack_kernel<<<dim3((unsigned)gx, 3), QK_THREADS, smem, st>>>(p);
    count_launches(1);
    return check_launch("qkv_pack");
}

}  // namespace pdsc
