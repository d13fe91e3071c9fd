// Throughput of packed fp32 FMA (fma.rn.f32x2 -> FFMA2) against scalar FFMA on sm_100a.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_bench.bin ffma2_bench.cu
#include <cuda_runtime.h>
#include <stdio.h>
__device__ __forceinline__ void ffma2(float2& d, float2 a, float2 b) {
    asm volatile("{\n\t.reg .b64 ra, rb, rd;\n\t"
        "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\tmov.b64 rd, {%0, %1};\n\t"
        "fma.rn.f32x2 rd, ra, rb, rd;\n\t"
        "mov.b64 {%0, %1}, rd;\n\t}"
        : "+f"(d.x), "+f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
}
template <int PACKED>
__global__ void __launch_bounds__(256) k(float* out, int iters, float s) {
    float2 acc[8];
    for (int i = 0; i < 8; ++i) acc[i] = make_float2(threadIdx.x * 1e-3f + i, i * 0.5f);
    const float2 a = make_float2(s, s * 0.999f), b = make_float2(1.0001f, 0.9999f);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (PACKED) ffma2(acc[i], a, b);
            else { acc[i].x = fmaf(a.x, b.x, acc[i].x); acc[i].y = fmaf(a.y, b.y, acc[i].y); }
        }
    }
    float r = 0.f;
    for (int i = 0; i < 8; ++i) r += acc[i].x + acc[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
int main() {
    float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 200000;
    for (int p = 0; p < 2; ++p) {
        for (int rep = 0; rep < 2; ++rep) {
            cudaEventRecord(e0);
            if (p) k<1><<<148 * 8, 256>>>(d, iters, 1e-6f); else k<0><<<148 * 8, 256>>>(d, iters, 1e-6f);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
        }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double flops = 2.0 * 16 * iters * 148.0 * 8 * 256;
        printf("%s: %.2f ms, %.1f TFLOP/s\n", p ? "FFMA2 (fma.rn.f32x2)" : "FFMA scalar", ms, flops / ms / 1e9);
    }
    return 0;
}
