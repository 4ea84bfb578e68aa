// Micro-benchmark: tcgen05.mma kind::f16 issue/execution rate for M=128 and N in {32,64,128,256}, A from TMEM
// (.ts) or shared memory (.ss), issued by one elected lane in straight-line runs of 8.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../pointdsc_b200/csrc -o mma_rate mma_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "sm100.cuh"
using namespace pdsc::sm100;

template <int N, bool TS>
__global__ void __launch_bounds__(128, 1) k(long long* cycles, int runs) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    __shared__ uint32_t tmem_base;
    __shared__ uint64_t bar;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    if (warp == 0) tmem_alloc(&tmem_base, 512);
    for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm)[i] = 0;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 1) {
        const uint64_t da = umma_desc_sw128_kmajor(smem_u32(sm));
        const uint64_t db = umma_desc_sw128_kmajor(smem_u32(sm + 32768));
        constexpr uint32_t idesc = umma_idesc_f16(128, N);
        long long t0 = clock64();
        for (int r = 0; r < runs; ++r) {
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {
                const uint64_t b = db + (uint64_t)((kk & 3) * 2);
                if (TS) umma_ts_elect(0u, 256u + (uint32_t)(kk & 3) * 8u, b, idesc, 1u);
                else umma_ss_elect(0u, da + (uint64_t)((kk & 3) * 2), b, idesc, 1u);
            }
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

template <int N, bool TS>
void run(const char* name) {
    long long* cyc;
    cudaMalloc(&cyc, 16);
    cudaFuncSetAttribute(k<N, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 66560);
    const int runs = 200;
    k<N, TS><<<1, 128, 66560>>>(cyc, 10);
    k<N, TS><<<1, 128, 66560>>>(cyc, runs);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[2];
    cudaMemcpy(h, cyc, 16, cudaMemcpyDeviceToHost);
    printf("%-24s issue %6.1f clk/MMA, issue+drain %6.1f clk/MMA  (%s)\n", name, (double)h[0] / (runs * 8), (double)h[1] / (runs * 8),
           cudaGetErrorString(e));
    cudaFree(cyc);
}

int main() {
    run<32, true>("M128 N32  K16 ts");
    run<64, true>("M128 N64  K16 ts");
    run<128, true>("M128 N128 K16 ts");
    run<256, true>("M128 N256 K16 ts");
    run<32, false>("M128 N32  K16 ss");
    run<64, false>("M128 N64  K16 ss");
    run<128, false>("M128 N128 K16 ss");
    run<256, false>("M128 N256 K16 ss");
    return 0;
}
