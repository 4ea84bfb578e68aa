// fp32 -> fp16 hi/lo operand packing for the channel-last 1x1-conv GEMMs.
// The PointCN / projection / fc_message layers (models/PointDSC.py:12-23,54-62) stay library GEMMs
// (torch.mm -> cuBLAS), but run on the tensor cores at fp32-grade accuracy: every activation x is
// split as x = hi + lo (two fp16, 22 mantissa bits) and a row is laid out as
//     [ hi(K) | hi(K) | lo(K) | 1 1 0 .. ]            (Kp = round_up(3K+2, 8) halves)
// so that one fp16 GEMM with fp32 accumulation against [W_hi; W_lo; W_hi; b_hi; b_lo; 0] yields
// x W + b with the hi*hi + hi*lo + lo*hi terms.  This kernel fuses the elementwise work around a
// GEMM: optional residual add, optional ReLU, optional fp32 copy of the activated value.
#include "common.cuh"
#include "sm100.cuh"

namespace pdsc {

using namespace sm100;

__global__ void __launch_bounds__(256) split16_rows_kernel(const float* __restrict__ x, const float* __restrict__ add,
                                                           float* __restrict__ y_out, uint8_t* __restrict__ out16,
                                                           long long rows, int K, int Kp, int relu) {
    const int per_row = K >> 3;   // 8-channel chunks per row
    const long long task = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (task >= rows * per_row) return;
    const long long r = task / per_row;
    const int c8 = (int)(task - r * per_row);
    const float4* px = reinterpret_cast<const float4*>(x + r * K + c8 * 8);
    float4 a0 = px[0], a1 = px[1];
    if (add) {
        const float4* pa = reinterpret_cast<const float4*>(add + r * K + c8 * 8);
        const float4 b0 = pa[0], b1 = pa[1];
        a0.x += b0.x; a0.y += b0.y; a0.z += b0.z; a0.w += b0.w;
        a1.x += b1.x; a1.y += b1.y; a1.z += b1.z; a1.w += b1.w;
    }
    if (relu) {
        a0.x = fmaxf(a0.x, 0.f); a0.y = fmaxf(a0.y, 0.f); a0.z = fmaxf(a0.z, 0.f); a0.w = fmaxf(a0.w, 0.f);
        a1.x = fmaxf(a1.x, 0.f); a1.y = fmaxf(a1.y, 0.f); a1.z = fmaxf(a1.z, 0.f); a1.w = fmaxf(a1.w, 0.f);
    }
    if (y_out) {
        float4* py = reinterpret_cast<float4*>(y_out + r * K + c8 * 8);
        py[0] = a0;
        py[1] = a1;
    }
    const float v[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
    uint32_t h[4], l[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        h[i] = pack_half2(v[2 * i], v[2 * i + 1]);
        const float2 back = unpack_half2(h[i]);
        l[i] = pack_half2(v[2 * i] - back.x, v[2 * i + 1] - back.y);
    }
    uint8_t* row = out16 + (size_t)r * Kp * 2;
    const uint4 hi = make_uint4(h[0], h[1], h[2], h[3]), lo = make_uint4(l[0], l[1], l[2], l[3]);
    *reinterpret_cast<uint4*>(row + (size_t)(c8 * 8) * 2) = hi;
    *reinterpret_cast<uint4*>(row + (size_t)(K + c8 * 8) * 2) = hi;
    *reinterpret_cast<uint4*>(row + (size_t)(2 * K + c8 * 8) * 2) = lo;
    if (c8 == 0) {   // bias columns: two ones, then zero padding up to Kp
        const int tail = Kp - 3 * K;   // multiple of ... at least 2
        __half* t = reinterpret_cast<__half*>(row) + 3 * K;
        for (int i = 0; i < tail; ++i) t[i] = __float2half(i < 2 ? 1.0f : 0.0f);
    }
}

int launch_split16_rows(const float* x, const float* add, float* y_out, void* out16, long long rows, int K, int Kp,
                        int relu, cudaStream_t st) {
    const long long tasks = rows * (K >> 3);
    const int blocks = (int)((tasks + 255) / 256);
    split16_rows_kernel<<<blocks, 256, 0, st>>>(x, add, y_out, reinterpret_cast<uint8_t*>(out16), rows, K, Kp, relu);
    count_launches(1);
    return check_launch("split16_rows");
}

}  // namespace pdsc
