// Seed kNN, stage 1 on the tensor cores: D[b][s][j] = 2 - 2 <f_seed(s), f_j>  (self -> +inf) for the S seed rows of every
// pair (knn, models/common.py:48-69 called at models/PointDSC.py:251-252, restricted to the rows the caller keeps).
// The fp32 CUDA-core Gram (knn_dist_kernel, seeds.cu) is bound by shared-memory bandwidth at ~43 % of the FMA peak; here
// the same numbers come from tcgen05 with fp16 hi/lo split operands (hi*lo + lo*hi + hi*hi, fp32 accumulation in TMEM:
// fp32-grade accuracy, the small cross terms are accumulated first so that the truncating accumulator loses least):
//   * knn_pack_seeds_kernel gathers the seed rows of a pair, 256 at a time, into the shared-memory image of a K-major
//     tcgen05 B operand (fp16 hi image, then lo; 64-channel blocks of 128-byte rows, 128B swizzle);
//   * knn_gram_tc_kernel: a work item is (pair, 256 seeds, 128 candidate rows).  A CTA takes a contiguous range of
//     items, so the 128 KB seed image is (re)loaded about twice per CTA; candidate rows are staged with cp.async one
//     item ahead, split into the A operand in TMEM by two threads per row, multiplied by 24 MMAs of 128 x 256 x 16, and
//     the 128 x 256 accumulator leaves transposed: for one seed the 32 lanes of a warp write 32 consecutive floats.
#include "common.cuh"
#include "sm100.cuh"

#include <math.h>

namespace pdsc {

using namespace sm100;

constexpr int KT_THREADS = 256;
constexpr int KT_SEEDS = 256;                                   // seeds per B image
constexpr uint32_t KT_BLK = KT_SEEDS * 128u;                    // one 64-channel block: 256 rows x 128 B
constexpr uint32_t KT_IMG = 4u * KT_BLK;                        // hi blk0 | hi blk1 | lo blk0 | lo blk1 = 128 KB
constexpr uint32_t KT_PITCH = 528;                              // staged candidate rows: 512 B + 16 (conflict-free LDS.128 per row)
constexpr uint32_t KT_STAGE = 128 * KT_PITCH;
constexpr uint32_t KT_D = 128;                                  // TMEM columns: A hi [0,64), A lo [64,128), D [128,384)

size_t knn_tc_image_bytes(int B, int S) { return (size_t)B * ((S + KT_SEEDS - 1) / KT_SEEDS) * KT_IMG; }

__device__ __forceinline__ void kt_split_pair(float a, float b, uint32_t& hi, uint32_t& lo) {
    hi = pack_half2(a, b);
    const float2 back = unpack_half2(hi);
    lo = pack_half2(a - back.x, b - back.y);
}

// grid (halves per pair, B), 256 threads: thread r packs seed row h*256 + r (zeros beyond S)
__global__ void __launch_bounds__(KT_SEEDS) knn_pack_seeds_kernel(const float* __restrict__ feat, const int64_t* __restrict__ seeds,
                                                                  int N, int S, uint8_t* __restrict__ bimg) {
    const int h = blockIdx.x, b = blockIdx.y, r = threadIdx.x;
    const int s = h * KT_SEEDS + r;
    uint8_t* img = bimg + ((size_t)b * gridDim.x + h) * KT_IMG;
    const float4* row = s < S ? reinterpret_cast<const float4*>(feat + ((size_t)b * N + (size_t)seeds[(size_t)b * S + s]) * 128) : nullptr;
#pragma unroll 4
    for (int chunk = 0; chunk < 16; ++chunk) {                  // 8 channels = 16 bytes of fp16 per chunk
        uint4 hi = make_uint4(0u, 0u, 0u, 0u), lo = hi;
        if (row) {
            const float4 v0 = __ldg(row + 2 * chunk), v1 = __ldg(row + 2 * chunk + 1);
            kt_split_pair(v0.x, v0.y, hi.x, lo.x);
            kt_split_pair(v0.z, v0.w, hi.y, lo.y);
            kt_split_pair(v1.x, v1.y, hi.z, lo.z);
            kt_split_pair(v1.z, v1.w, hi.w, lo.w);
        }
        const uint32_t off = (uint32_t)(chunk >> 3) * KT_BLK + (uint32_t)r * 128u + ((((uint32_t)chunk & 7u) ^ ((uint32_t)r & 7u)) << 4);
        *reinterpret_cast<uint4*>(img + off) = hi;
        *reinterpret_cast<uint4*>(img + 2 * KT_BLK + off) = lo;
    }
}

struct KtParams {
    const float* feat; const int64_t* seeds; const uint8_t* bimg; float* D;
    int N, S, nh, NT;          // halves (seed images) per pair, 128-row candidate tiles per pair
    long long items;           // B * nh * NT
};

__device__ __forceinline__ void kt_stage_rows(uint8_t* dst, const float* src, int nrows) {
    const int total = nrows << 5;                               // 32 chunks of 16 bytes per row
#pragma unroll 4
    for (int idx = threadIdx.x; idx < total; idx += KT_THREADS) {
        const int r = idx >> 5, c = idx & 31;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst + (uint32_t)r * KT_PITCH + (uint32_t)c * 16u)),
                     "l"(src + (size_t)r * 128 + c * 4)
                     : "memory");
    }
}

__global__ void __launch_bounds__(KT_THREADS, 1) knn_gram_tc_kernel(const KtParams p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* bsm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* xs = bsm + KT_IMG;
    __shared__ int seed_idx[KT_SEEDS];
    __shared__ uint64_t w_bar, mma_bar;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int rt = tid & 127, half = tid >> 7;                  // row of the tile, which 32-column chunks of it

    if (tid == 0) {
        mbar_init(&w_bar, 1);
        mbar_init(&mma_bar, 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(&tmem_base_s, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;
    const uint32_t lane_addr = (uint32_t)((warp & 3) * 32) << 16;

    const long long i0 = p.items * blockIdx.x / gridDim.x, i1 = p.items * (blockIdx.x + 1) / gridDim.x;
    auto stage_item = [&](long long item) {
        const long long g = item / p.NT;
        const int t = (int)(item - g * p.NT), b = (int)(g / p.nh);
        kt_stage_rows(xs, p.feat + ((size_t)b * p.N + (size_t)t * 128) * 128, min(128, p.N - t * 128));
    };
    if (i0 < i1) stage_item(i0);
    asm volatile("cp.async.commit_group;" ::: "memory");

    long long cur_g = -1;
    uint32_t w_phase = 0, mma_phase = 0;
    for (long long item = i0; item < i1; ++item) {
        const long long g = item / p.NT;
        const int t = (int)(item - g * p.NT), b = (int)(g / p.nh), h = (int)(g - (long long)b * p.nh);
        const bool new_g = g != cur_g;
        if (new_g) {
            __syncthreads();           // the previous item's epilogue (seed_idx) is over; its MMAs were waited for
            if (tid == 0) {
                mbar_arrive_expect_tx(&w_bar, KT_IMG);
                const uint8_t* src = p.bimg + (size_t)g * KT_IMG;
#pragma unroll
                for (uint32_t off = 0; off < KT_IMG; off += 32768u) bulk_g2s(bsm + off, src + off, 32768u, &w_bar);
            }
            const int s = h * KT_SEEDS + tid;
            seed_idx[tid] = s < p.S ? (int)p.seeds[(size_t)b * p.S + s] : -1;
            cur_g = g;
        }
        // ---- candidate rows (staged) -> A operand (hi, lo) in TMEM ----
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();
        {
            const float4* xr = reinterpret_cast<const float4*>(xs + (uint32_t)rt * KT_PITCH);
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
                const int c = 2 * cc + half;
                uint32_t hi[16], lo[16];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float4 v = xr[c * 8 + i];
                    kt_split_pair(v.x, v.y, hi[2 * i], lo[2 * i]);
                    kt_split_pair(v.z, v.w, hi[2 * i + 1], lo[2 * i + 1]);
                }
                tmem_st16(tmem + lane_addr + (uint32_t)c * 16u, hi);
                tmem_st16(tmem + lane_addr + 64u + (uint32_t)c * 16u, lo);
            }
        }
        tmem_st_wait();
        tc_fence_before();
        __syncthreads();               // A operand complete, the stage has been read, D of the previous item has been read
        if (new_g) { mbar_wait(&w_bar, w_phase); w_phase ^= 1; }
        if (warp == 0) {
            tc_fence_after();
            constexpr uint32_t idesc = umma_idesc_f16(128, KT_SEEDS);
            const uint64_t d0 = umma_desc_sw128_kmajor(smem_u32(bsm));
            const uint32_t dlo = __shfl_sync(0xffffffffu, (uint32_t)d0, 0), dhi = __shfl_sync(0xffffffffu, (uint32_t)(d0 >> 32), 0);
            const uint64_t wd = ((uint64_t)dhi << 32) | dlo;
            uint32_t acc = 0;
#pragma unroll
            for (int term = 0; term < 3; ++term) {             // A hi x B lo, A lo x B hi, A hi x B hi
                const uint32_t a0 = tmem + (term == 1 ? 64u : 0u);
                const uint32_t b0 = term == 0 ? 2u * KT_BLK : 0u;
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    const uint64_t db = wd + (uint64_t)((b0 + (uint32_t)(kk >> 2) * KT_BLK + (uint32_t)(kk & 3) * 32u) >> 4);
                    umma_ts_elect(tmem + KT_D, a0 + (uint32_t)kk * 8u, db, idesc, acc);
                    acc = 1;
                }
            }
            umma_commit_elect(&mma_bar);
        }
        if (item + 1 < i1) stage_item(item + 1);               // the stage is free: next item's rows fly during the MMAs / epilogue
        asm volatile("cp.async.commit_group;" ::: "memory");
        mbar_wait(&mma_bar, mma_phase);
        mma_phase ^= 1;
        tc_fence_after();
        // ---- epilogue: 2 - 2 dot, self -> +inf, transposed write-out ----
        {
            const int cand = t * 128 + rt;
            const bool live = cand < p.N;
            float* dcol = p.D + ((size_t)b * p.S + (size_t)h * KT_SEEDS) * p.N + cand;
            const int s_left = p.S - h * KT_SEEDS;              // seeds of this image that exist
#pragma unroll 1
            for (int cc = 0; cc < 4; ++cc) {
                const int c = 2 * cc + half;
                if (c * 32 >= s_left) break;                    // warp-uniform
                uint32_t r[32];
                tmem_ld32(tmem + lane_addr + KT_D + (uint32_t)c * 32u, r);
                tmem_ld_wait();
                if (live) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const int sl = c * 32 + i;
                        if (sl < s_left) {
                            const float v = seed_idx[sl] == cand ? INFINITY : __fsub_rn(2.0f, __fmul_rn(2.0f, __uint_as_float(r[i])));   // common.py:58-60
                            __stcs(dcol + (size_t)sl * p.N, v);
                        }
                    }
                }
            }
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
}

int launch_knn_gram_tc(const float* normed, const int64_t* seeds, int B, int N, int S, float* D, void* img, cudaStream_t st) {
    const int nh = (S + KT_SEEDS - 1) / KT_SEEDS;
    knn_pack_seeds_kernel<<<dim3(nh, B), KT_SEEDS, 0, st>>>(normed, seeds, N, S, reinterpret_cast<uint8_t*>(img));
    KtParams p;
    p.feat = normed; p.seeds = seeds; p.bimg = reinterpret_cast<const uint8_t*>(img); p.D = D;
    p.N = N; p.S = S; p.nh = nh; p.NT = (N + 127) / 128;
    p.items = (long long)B * nh * p.NT;
    const int smem = (int)(KT_IMG + KT_STAGE + 1024);
    opt_in_smem<knn_gram_tc_kernel>(smem);
    long long grid = num_sms();
    if (grid > p.items) grid = p.items;
    knn_gram_tc_kernel<<<(int)grid, KT_THREADS, smem, st>>>(p);
    count_launches(2);
    return check_launch("seed_knn (tensor-core Gram)");
}

}  // namespace pdsc
