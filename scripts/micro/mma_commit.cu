// Micro-benchmark: what the bookkeeping between groups of tcgen05.mma costs the issuing warp / the tensor pipe.
// Groups of G MMAs (M128 N64 K16, A from TMEM) separated by: nothing | 1 commit | 3 commits | commit + try_wait on a
// completed barrier + tcgen05.fence | the same plus a clock read and a global store (the trace hook).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../pointdsc_b200/csrc -o mma_commit mma_commit.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "sm100.cuh"
using namespace pdsc::sm100;

template <int MODE, int G>
__global__ void __launch_bounds__(128, 1) k(long long* cycles, int runs) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    __shared__ uint32_t tmem_base;
    __shared__ uint64_t bar, junk[3], done;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1); mbar_init(&done, 1);
        for (int i = 0; i < 3; ++i) mbar_init(&junk[i], 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(&tmem_base, 512);
    for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm)[i] = 0;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (threadIdx.x == 0) mbar_arrive(&done);       // `done` has completed phase 0: waits on parity 0 return at once
    __syncthreads();
    if (warp == 1) {
        const uint64_t db = umma_desc_sw128_kmajor(smem_u32(sm + 32768));
        constexpr uint32_t idesc = umma_idesc_f16(128, 64);
        long long t0 = clock64();
        for (int r = 0; r < runs; ++r) {
#pragma unroll
            for (int kk = 0; kk < G; ++kk)
                umma_ts_elect(0u, 256u + (uint32_t)(kk & 3) * 8u, db + (uint64_t)((kk & 3) * 2), idesc, 1u);
            if (MODE >= 1) umma_commit_elect(&junk[0]);
            if (MODE == 2) { umma_commit_elect(&junk[1]); umma_commit_elect(&junk[2]); }
            if (MODE >= 3) { mbar_wait_uniform(&done, 0); tc_fence_after(); }
            if (MODE >= 4) { if (threadIdx.x == 32) cycles[8 + (r & 7)] = clock64(); }
        }
        long long t1 = clock64();
        umma_commit_elect(&bar);
        mbar_wait_uniform(&bar, 0);
        long long t2 = clock64();
        if (threadIdx.x == 32) { cycles[0] = t1 - t0; cycles[1] = t2 - t0; }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(0u, 512);
}

template <int MODE, int G>
void run(const char* name) {
    long long* cyc;
    cudaMalloc(&cyc, 256);
    cudaFuncSetAttribute(k<MODE, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, 66560);
    const int runs = 200;
    k<MODE, G><<<1, 128, 66560>>>(cyc, 10);
    k<MODE, G><<<1, 128, 66560>>>(cyc, runs);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[2];
    cudaMemcpy(h, cyc, 16, cudaMemcpyDeviceToHost);
    printf("%-44s G=%2d: issue %7.1f clk/group, issue+drain %7.1f clk/group = %5.1f clk/MMA (%s)\n", name, G, (double)h[0] / runs,
           (double)h[1] / runs, (double)h[1] / (runs * G), cudaGetErrorString(e));
    cudaFree(cyc);
}

int main() {
    run<0, 8>("no bookkeeping");
    run<1, 8>("1 commit");
    run<2, 8>("3 commits");
    run<3, 8>("commit + wait(done) + fence");
    run<4, 8>("commit + wait + fence + clock/store");
    run<0, 24>("no bookkeeping");
    run<1, 24>("1 commit");
    run<2, 24>("3 commits");
    run<3, 24>("commit + wait(done) + fence");
    run<4, 24>("commit + wait + fence + clock/store");
    run<0, 1>("no bookkeeping");
    run<1, 1>("1 commit");
    return 0;
}
