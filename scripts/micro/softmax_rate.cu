// Micro-benchmark: the per-tile softmax arithmetic of the attention kernel in isolation (no MMA, no
// TMEM): cycles per 128x128 tile per SM for NWG x 128 threads.  Variants select the instruction mix.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o softmax_rate softmax_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
__device__ __forceinline__ float fast_sqrt(float x) { float y; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float fast_rsqrt(float x) { float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float fast_exp2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float fma_sat(float a, float b, float c) { float d; asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
__device__ __forceinline__ uint32_t pack_half2(float a, float b) { __half2 h = __floats2half2_rn(a, b); return *reinterpret_cast<uint32_t*>(&h); }
__device__ __forceinline__ float2 unpack_half2(uint32_t u) { __half2 h = *reinterpret_cast<__half2*>(&u); return __half22float2(h); }

// VAR bit 0: compat part, bit 1: fma+max part, bit 2: exp part, bit 3: hi/lo split
template <int NP, int VAR>
__global__ void __launch_bounds__(512, 1) k(const float4* __restrict__ gpts, uint32_t* out, int tiles, long long* cycles) {
    __shared__ float4 pts[64][3];
    for (int i = threadIdx.x; i < 192; i += blockDim.x) pts[i / 3][i % 3] = gpts[i];
    __syncthreads();
    const int wg = threadIdx.x >> 7;
    const int kp0 = wg * NP;
    float2 qp[6];
    for (int i = 0; i < 6; ++i) qp[i] = make_float2(0.1f * threadIdx.x + i, 0.1f * threadIdx.x + i);
    float m = 0.f, l = 0.f;
    uint32_t sink = 0;
    float sfake = 0.01f * threadIdx.x;
    long long t0 = clock64();
    for (int j = 0; j < tiles; ++j) {
        float2 e[NP];
#pragma unroll
        for (int i = 0; i < 6; ++i) qp[i] = __fadd2_rn(qp[i], make_float2(sfake, sfake));   // defeat hoisting
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            if (VAR & 1) {
                const float4 pa = pts[kp0 + i][0], pb = pts[kp0 + i][1], pc = pts[kp0 + i][2];
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
                e[i] = make_float2(fma_sat(-c.x, c.x, 1.0f), fma_sat(-c.y, c.y, 1.0f));
            } else {
                e[i] = make_float2(sfake + i, sfake - i);
            }
        }
        float emax = -1e30f;
        if (VAR & 2) {
            const float2 neg_m = make_float2(-m, -m);
#pragma unroll
            for (int i = 0; i < NP; ++i) {
                e[i] = __ffma2_rn(e[i], make_float2(sfake + j, sfake - i), neg_m);
                emax = fmaxf(emax, fmaxf(e[i].x, e[i].y));
            }
            if (emax > 1e20f) m += 1.f;
        }
        float2 psum = make_float2(0.f, 0.f);
        if (VAR & 4) {
#pragma unroll
            for (int i = 0; i < NP; ++i) {
                const float2 p = make_float2(fast_exp2(e[i].x), fast_exp2(e[i].y));
                psum = __fadd2_rn(psum, p);
                uint32_t hi = pack_half2(p.x, p.y);
                sink ^= hi;
                if (VAR & 8) {
                    const float2 back = unpack_half2(hi);
                    const float2 rem = __fadd2_rn(p, make_float2(-back.x, -back.y));
                    sink += pack_half2(rem.x, rem.y);
                }
            }
        } else {
#pragma unroll
            for (int i = 0; i < NP; ++i) sink ^= __float_as_uint(e[i].x) + __float_as_uint(e[i].y);
        }
        l += psum.x + psum.y;
        sfake += 1e-3f;
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = sink + __float_as_uint(l);
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}
template <int NP, int VAR>
void run(const char* name, int nthreads) {
    float4* pts; uint32_t* out; long long* cyc;
    cudaMalloc(&pts, 192 * 16); cudaMemset(pts, 0, 192 * 16);
    cudaMalloc(&out, 4 * 148 * 1024); cudaMalloc(&cyc, 8);
    const int tiles = 2000;
    k<NP, VAR><<<148, nthreads>>>(pts, out, 10, cyc);
    k<NP, VAR><<<148, nthreads>>>(pts, out, tiles, cyc);
    cudaDeviceSynchronize();
    long long h;
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-28s threads %4d: %7.0f cycles per 128x128 tile\n", name, nthreads, (double)h / tiles);
}
int main() {
    run<16, 15>("all (NWG=4, 32 keys/thr)", 512);
    run<32, 15>("all (NWG=2, 64 keys/thr)", 256);
    run<16, 1>("compat only", 512);
    run<16, 2>("fma+max only", 512);
    run<16, 4>("exp only (no lo split)", 512);
    run<16, 12>("exp + hi/lo split", 512);
    run<16, 7>("all but lo split", 512);
    printf("cuda error: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
