// Hardware check of the tcgen05.ld/st .16x32bx2 shape: which (lane, column) each thread sees, and
// whether the 16-lane window may start at lane 16 of the warp's quarter.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../pointdsc_b200/csrc -o tmem_shapes tmem_shapes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "sm100.cuh"
using namespace pdsc::sm100;

__device__ __forceinline__ void ld_16x32bx2_x4(uint32_t taddr, uint32_t (&r)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.16x32bx2.x4.b32 {%0, %1, %2, %3}, [%4], 8;"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void st_16x32bx2_x4(uint32_t taddr, const uint32_t (&r)[4]) {
    asm volatile("tcgen05.st.sync.aligned.16x32bx2.x4.b32 [%0], 8, {%1, %2, %3, %4};"
                 ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]) : "memory");
}

__global__ void k(uint32_t* out) {
    __shared__ uint32_t tmem_base;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) tmem_alloc(&tmem_base, 32);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base;
    const uint32_t quarter = (uint32_t)(warp & 3) * 32u;
    // pattern: value = (lane << 8) | column, written with the plain 32x32b shape
    {
        uint32_t r[16];
        for (int i = 0; i < 16; ++i) r[i] = ((quarter + lane) << 8) | i;
        tmem_st16(tmem + (quarter << 16), r);
        tmem_st_wait();
    }
    __syncthreads();
    // read with 16x32bx2 at lane offsets 0 and 16: 4 regs, second half 8 columns further
    for (int half = 0; half < 2; ++half) {
        uint32_t r[4];
        ld_16x32bx2_x4(tmem + ((quarter + 16u * half) << 16), r);
        tmem_ld_wait();
        for (int i = 0; i < 4; ++i) out[((warp * 2 + half) * 32 + lane) * 4 + i] = r[i];
    }
    __syncthreads();
    // write with 16x32bx2 (lane offset 16*half), read back with 32x32b
    for (int half = 0; half < 2; ++half) {
        uint32_t r[4];
        for (int i = 0; i < 4; ++i) r[i] = 0xA0000000u | (half << 24) | (lane << 8) | i;
        st_16x32bx2_x4(tmem + ((quarter + 16u * half) << 16), r);
    }
    tmem_st_wait();
    __syncthreads();
    {
        uint32_t r[16];
        tmem_ld16(tmem + (quarter << 16), r);
        tmem_ld_wait();
        for (int i = 0; i < 16; ++i) out[4 * 2 * 32 * 4 + (warp * 32 + lane) * 16 + i] = r[i];
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 32);
}

int main() {
    uint32_t* d;
    const int n = 4 * 2 * 32 * 4 + 4 * 32 * 16;
    cudaMalloc(&d, n * 4);
    cudaMemset(d, 0xff, n * 4);
    k<<<1, 128>>>(d);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    static uint32_t h[4 * 2 * 32 * 4 + 4 * 32 * 16];
    cudaMemcpy(h, d, n * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int w = 0; w < 4; ++w)
        for (int half = 0; half < 2; ++half)
            for (int t = 0; t < 32; ++t)
                for (int i = 0; i < 4; ++i) {
                    const uint32_t got = h[((w * 2 + half) * 32 + t) * 4 + i];
                    const uint32_t exp_lane = w * 32 + 16 * half + (t & 15), exp_col = (t >> 4) * 8 + i;
                    if (got != ((exp_lane << 8) | exp_col)) {
                        if (bad < 8) printf("ld mismatch w%d half%d t%d i%d: got lane %u col %u, expected lane %u col %u\n", w, half, t, i, got >> 8, got & 255, exp_lane, exp_col);
                        ++bad;
                    }
                }
    printf("16x32bx2 load check: %d mismatches\n", bad);
    bad = 0;
    const uint32_t* h2 = h + 4 * 2 * 32 * 4;
    for (int w = 0; w < 4; ++w)
        for (int t = 0; t < 32; ++t)
            for (int c = 0; c < 16; ++c) {
                const uint32_t got = h2[(w * 32 + t) * 16 + c];
                // lane t of the quarter, column c: written by half = t/16, thread (t%16) + 16*(c/8), reg c%8 (only c%8 < 4 written)
                if ((c & 7) < 4) {
                    const uint32_t exp = 0xA0000000u | ((uint32_t)(t >> 4) << 24) | ((uint32_t)((t & 15) + 16 * (c >> 3)) << 8) | (c & 7);
                    if (got != exp) {
                        if (bad < 8) printf("st mismatch w%d lane%d col%d: got %08x expected %08x\n", w, t, c, got, exp);
                        ++bad;
                    }
                }
            }
    printf("16x32bx2 store check: %d mismatches\n", bad);
    return 0;
}
