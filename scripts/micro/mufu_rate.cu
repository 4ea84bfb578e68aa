// Micro-benchmark: per-SM throughput of the XU (MUFU) ops the attention softmax uses.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mufu_rate mufu_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int OP>
__global__ void k(float* out, int iters) {
    float x[8];
    for (int i = 0; i < 8; ++i) x[i] = 1.0f + threadIdx.x * 1e-3f + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (OP == 0) asm volatile("sqrt.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
            if (OP == 1) asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
            if (OP == 2) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
            if (OP == 3) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
            if (OP == 4) asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(x[i]));
            if (OP == 5) asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
        }
    }
    float s = 0;
    for (int i = 0; i < 8; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int OP>
void run(const char* name, int sms) {
    float* out;
    cudaMalloc(&out, sizeof(float) * sms * 1024);
    const int iters = 20000;
    k<OP><<<sms, 1024>>>(out, 10);
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    k<OP><<<sms, 1024>>>(out, iters);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    double ops = (double)sms * 1024 * 8 * iters;
    int clk;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("%-6s %.1f Gop/s total, %.2f ops/ns/SM  (at %d MHz nominal: %.1f ops/clk/SM)\n", name, ops / ms / 1e6,
           ops / ms / 1e6 / sms, clk / 1000, ops / ms / 1e6 / sms / (clk / 1e6));
    cudaFree(out);
}
int main() {
    int sms;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    run<0>("sqrt", sms); run<1>("rsqrt", sms); run<2>("ex2", sms); run<3>("rcp", sms); run<5>("lg2", sms); run<4>("ffma", sms);
    return 0;
}
