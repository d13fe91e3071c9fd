// Does ptxas for sm_100a accept the packed fp32 FMA, and what SASS does it emit?
#include <cuda_runtime.h>
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    float2 d;
    asm("{\n\t.reg .b64 ra, rb, rc, rd;\n\t"
        "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\tmov.b64 rc, {%6, %7};\n\t"
        "fma.rn.f32x2 rd, ra, rb, rc;\n\t"
        "mov.b64 {%0, %1}, rd;\n\t}"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
    return d;
}
__global__ void k(const float2* a, const float2* b, float2* c, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    float2 acc = make_float2(0.f, 0.f);
    for (int j = 0; j < n; ++j) acc = ffma2(a[i + j], b[i + j], acc);
    c[i] = acc;
}
