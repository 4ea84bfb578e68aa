// Micro-benchmark: cost of the broadcast shared-memory loads of key points vs the compat arithmetic.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ float fast_sqrt(float x) { float y; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float fma_sat(float a, float b, float c) { float d; asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }

// MODE 0: LDS.128 x3 per pair + arithmetic; 1: LDS only (xor sink); 2: arithmetic only (key data from registers, perturbed)
// 3: LDS.32 x12 only; 4: arithmetic with scalar (non-packed) FP32 ops, no LDS; 5: LDS.64 x6 only
template <int MODE>
__global__ void __launch_bounds__(1024, 1) k(const float4* __restrict__ gpts, uint32_t* out, int tiles, long long* cycles) {
    __shared__ float4 pts[2][64][3];
    for (int i = threadIdx.x; i < 384; i += blockDim.x) (&pts[0][0][0])[i] = gpts[i % 192];
    __syncthreads();
    const int NP = 16;
    const int wg = threadIdx.x >> 7;
    const int kp0 = wg * NP;
    float2 qp[6];
    for (int i = 0; i < 6; ++i) qp[i] = make_float2(0.1f * threadIdx.x + i, 0.1f * threadIdx.x + i);
    uint32_t sink = 0;
    float acc = 0.f;
    long long t0 = clock64();
    for (int j = 0; j < tiles; ++j) {
        const float4 (*P)[3] = &pts[j & 1][kp0];
        const float pert = 1e-3f * j;
#pragma unroll
        for (int i = 0; i < 6; ++i) qp[i] = __fadd2_rn(qp[i], make_float2(pert, pert));
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            float4 pa, pb, pc;
            if (MODE == 0 || MODE == 1) { pa = P[i][0]; pb = P[i][1]; pc = P[i][2]; }
            if (MODE == 3) {
                const volatile float* f = reinterpret_cast<const volatile float*>(&P[i][0]);
                pa = make_float4(f[0], f[1], f[2], f[3]); pb = make_float4(f[4], f[5], f[6], f[7]); pc = make_float4(f[8], f[9], f[10], f[11]);
            }
            if (MODE == 5) {
                const volatile float* f = reinterpret_cast<const volatile float*>(&P[i][0]);
                float2 a, b, c, d, e, g;
                asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(a.x), "=f"(a.y) : "r"((uint32_t)__cvta_generic_to_shared((const void*)(f + 0))));
                asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(b.x), "=f"(b.y) : "r"((uint32_t)__cvta_generic_to_shared((const void*)(f + 2))));
                asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(c.x), "=f"(c.y) : "r"((uint32_t)__cvta_generic_to_shared((const void*)(f + 4))));
                asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(d.x), "=f"(d.y) : "r"((uint32_t)__cvta_generic_to_shared((const void*)(f + 6))));
                asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(e.x), "=f"(e.y) : "r"((uint32_t)__cvta_generic_to_shared((const void*)(f + 8))));
                asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(g.x), "=f"(g.y) : "r"((uint32_t)__cvta_generic_to_shared((const void*)(f + 10))));
                pa = make_float4(a.x, a.y, b.x, b.y); pb = make_float4(c.x, c.y, d.x, d.y); pc = make_float4(e.x, e.y, g.x, g.y);
            }
            if (MODE == 2 || MODE == 4) {
                pa = make_float4(pert + i, pert - i, pert * 2 + i, pert * 3 - i);
                pb = make_float4(pert + 2 * i, pert - 2 * i, pert * 5 + i, pert * 7 - i);
                pc = make_float4(pert + 3 * i, pert - 3 * i, pert * 11 + i, pert * 13 - i);
            }
            if (MODE == 1 || MODE == 3 || MODE == 5) {
                sink ^= __float_as_uint(pa.x) ^ __float_as_uint(pa.y) ^ __float_as_uint(pa.z) ^ __float_as_uint(pa.w) ^
                        __float_as_uint(pb.x) ^ __float_as_uint(pb.y) ^ __float_as_uint(pb.z) ^ __float_as_uint(pb.w) ^
                        __float_as_uint(pc.x) ^ __float_as_uint(pc.y) ^ __float_as_uint(pc.z) ^ __float_as_uint(pc.w);
            } else if (MODE == 4) {
                float w2[2];
                const float kx[2][6] = {{pa.x, pa.z, pb.x, pb.z, pc.x, pc.z}, {pa.y, pa.w, pb.y, pb.w, pc.y, pc.w}};
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    float dx = qp[0].x - kx[h][0], dy = qp[1].x - kx[h][1], dz = qp[2].x - kx[h][2];
                    float ds = fast_sqrt(fmaf(dz, dz, fmaf(dy, dy, dx * dx)));
                    dx = qp[3].x - kx[h][3]; dy = qp[4].x - kx[h][4]; dz = qp[5].x - kx[h][5];
                    float dt = fast_sqrt(fmaf(dz, dz, fmaf(dy, dy, dx * dx)));
                    float c = ds - dt;
                    w2[h] = fma_sat(-c, c, 1.0f);
                }
                acc += w2[0] + w2[1];
            } else {
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
                acc += fma_sat(-c.x, c.x, 1.0f) + fma_sat(-c.y, c.y, 1.0f);
            }
        }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = sink + __float_as_uint(acc);
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}
template <int MODE>
void run(const char* name, int nthreads) {
    float4* pts; uint32_t* out; long long* cyc;
    cudaMalloc(&pts, 192 * 16); cudaMemset(pts, 0, 192 * 16);
    cudaMalloc(&out, 4 * 148 * 1024); cudaMalloc(&cyc, 8);
    const int tiles = 2000;
    k<MODE><<<148, nthreads>>>(pts, out, 10, cyc);
    k<MODE><<<148, nthreads>>>(pts, out, tiles, cyc);
    cudaDeviceSynchronize();
    long long h;
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-44s threads %4d: %7.0f cycles per 128x128 tile\n", name, nthreads, (double)h / tiles);
}
int main() {
    run<0>("LDS.128 x3 + packed compat math", 512);
    run<1>("LDS.128 x3 only", 512);
    run<3>("LDS.32 x12 only", 512);
    run<5>("LDS.64 x6 only", 512);
    run<2>("packed compat math only (f32x2)", 512);
    run<4>("scalar compat math only", 512);
    run<0>("LDS.128 x3 + packed compat math", 256);
    run<2>("packed compat math only (f32x2)", 256);
    printf("cuda error: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
