// Shared device/host helpers for the MAT-AS B200 kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define MATAS_OK 0
#define MATAS_ERR_ARG 1
#define MATAS_ERR_CUDA 2
#define MATAS_ERR_UNSUPPORTED 3
#define MATAS_ERR_LIMIT 4

void matas_set_error(const char* fmt, ...);

#define CUDA_TRY(expr)                                                                         \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            matas_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return MATAS_ERR_CUDA;                                                             \
        }                                                                                      \
    } while (0)

#define ARG_CHECK(cond, ...)                \
    do {                                    \
        if (!(cond)) {                      \
            matas_set_error(__VA_ARGS__);   \
            return MATAS_ERR_ARG;           \
        }                                   \
    } while (0)

// ---------------------------------------------------------------------------------------------
// Philox-4x32-10: the named-draw generator (spec: oracle/philox.py, DESIGN.md "Random streams")
// ---------------------------------------------------------------------------------------------
enum { S_RETRY = 1, S_JITTER = 2, S_RESET = 3, S_WORKER = 4, S_POLICY = 8 };

__host__ __device__ __forceinline__ uint4 philox4x32(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                     uint64_t seed) {
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}
__host__ __device__ __forceinline__ uint32_t rng_tag(uint32_t stream, uint32_t index) { return (stream << 16) | index; }
__host__ __device__ __forceinline__ double u01d(uint32_t x) { return (double)(x >> 8) * (1.0 / 16777216.0); }
__host__ __device__ __forceinline__ float u01f(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }
__host__ __device__ __forceinline__ uint32_t bounded(uint32_t x, uint32_t n) {
    return (uint32_t)(((uint64_t)x * n) >> 32);
}
__host__ __device__ __forceinline__ uint32_t lane_of(const uint4& v, int i) {
    return i == 0 ? v.x : (i == 1 ? v.y : (i == 2 ? v.z : v.w));
}

// ---------------------------------------------------------------------------------------------
// math
// ---------------------------------------------------------------------------------------------
// erf to 1.5e-7 absolute (Abramowitz & Stegun 7.1.26): erf(|x|) = 1 - (a1 t + ... + a5 t^5) e^{-x^2}, t = 1 / (1 + p |x|).
// The exact-erf GELU of the reference (nn.GELU()) is evaluated with it in the forward epilogues: its error, 0.5 |x| 1.5e-7
// per activation, is at the level of fp32 rounding of the surrounding GEMM and an order of magnitude cheaper than
// erff's two-branch polynomial, which was a fifth of the decode's per-stage epilogue time.
__device__ __forceinline__ float erf_as(float x) {
    const float ax = fabsf(x);
    const float t = __fdividef(1.0f, fmaf(0.3275911f, ax, 1.0f));
    float p = fmaf(1.061405429f, t, -1.453152027f);
    p = fmaf(p, t, 1.421413741f);
    p = fmaf(p, t, -0.284496736f);
    p = fmaf(p, t, 0.254829592f);
    const float e = __expf(-ax * ax);
    return copysignf(fmaf(-p * t, e, 1.0f), x);
}
__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erf_as(x * 0.70710678118654752f)); }

// 256-bit global store (sm_100: st.global.v8.f32 -> STG.E.256): a thread that owns 32 or more consecutive bytes of a
// row writes whole 32-byte sectors, where two 128-bit stores would each write half a sector of a different instruction.
// p must be 32-byte aligned.
__device__ __forceinline__ void st_global_256(float* p, float a, float b, float c, float d, float e, float f, float g, float h) {
    asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d), "f"(e), "f"(f), "f"(g),
                 "f"(h)
                 : "memory");
}

// reduce over the 16 lanes of a half-warp (xor 8,4,2,1 stays inside the half)
// Packed fp32 arithmetic (sm_100: fma.rn.f32x2 -> SASS FFMA2).  Same FLOP rate as two FFMAs but ONE issue slot
// (measured on B200, scripts/tc/ffma2_bench.cu: 73.7 vs 72.6 TFLOP/s), which is what issue-bound loops need.
__device__ __forceinline__ void fma2(float& dx, float& dy, float ax, float ay, float bx, float by) {     // d += a * b
    asm("{\n\t.reg .b64 ra, rb, rd;\n\t"
        "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\tmov.b64 rd, {%0, %1};\n\t"
        "fma.rn.f32x2 rd, ra, rb, rd;\n\t"
        "mov.b64 {%0, %1}, rd;\n\t}"
        : "+f"(dx), "+f"(dy) : "f"(ax), "f"(ay), "f"(bx), "f"(by));
}
__device__ __forceinline__ void mul2(float& dx, float& dy, float s) {                                    // d *= s
    asm("{\n\t.reg .b64 rs, rd;\n\t"
        "mov.b64 rs, {%2, %2};\n\tmov.b64 rd, {%0, %1};\n\t"
        "mul.rn.f32x2 rd, rd, rs;\n\t"
        "mov.b64 {%0, %1}, rd;\n\t}"
        : "+f"(dx), "+f"(dy) : "f"(s));
}
__device__ __forceinline__ void fma2c(float& rx, float& ry, float ax, float ay, float bx, float by, float c) {   // r = a * b + (c, c)
    asm("{\n\t.reg .b64 ra, rb, rc;\n\t"
        "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\tmov.b64 rc, {%6, %6};\n\t"
        "fma.rn.f32x2 rc, ra, rb, rc;\n\t"
        "mov.b64 {%0, %1}, rc;\n\t}"
        : "=f"(rx), "=f"(ry) : "f"(ax), "f"(ay), "f"(bx), "f"(by), "f"(c));
}
__device__ __forceinline__ void mul2v(float& rx, float& ry, float ax, float ay, float bx, float by) {             // r = a * b
    asm("{\n\t.reg .b64 ra, rb;\n\t"
        "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\t"
        "mul.rn.f32x2 ra, ra, rb;\n\t"
        "mov.b64 {%0, %1}, ra;\n\t}"
        : "=f"(rx), "=f"(ry) : "f"(ax), "f"(ay), "f"(bx), "f"(by));
}
// Exact-erf GELU of two values at once, for the issue-bound epilogues (a [128 x 64] GELU stage on 16 warps is bound by the
// SM's 16 special-function lanes and its issue slots): erf(z) = 1 - 1 / P(z)^16, P = 1 + a1 z + ... + a6 z^6 (Abramowitz &
// Stegun 7.1.28, |err| <= 3e-7) needs ONE special-function op per value (the reciprocal) where 7.1.26 needs two, and its
// polynomial and squarings run as packed fp32x2 instructions.  gelu(x) = max(x, 0) - |x| / 2 . P(|x| / sqrt 2)^-16; measured
// against math.erf in fp32: 9e-7 absolute on the GELU output (the squarings amplify rounding 16-fold).
__device__ __forceinline__ void gelu_erf_x2(float& x0, float& x1) {
    const float a0 = fabsf(x0), a1 = fabsf(x1);
    const float z0 = a0 * 0.70710678118654752f, z1 = a1 * 0.70710678118654752f;
    float p0, p1;
    fma2c(p0, p1, 0.0000430638f, 0.0000430638f, z0, z1, 0.0002765672f);
    fma2c(p0, p1, p0, p1, z0, z1, 0.0001520143f);
    fma2c(p0, p1, p0, p1, z0, z1, 0.0092705272f);
    fma2c(p0, p1, p0, p1, z0, z1, 0.0422820123f);
    fma2c(p0, p1, p0, p1, z0, z1, 0.0705230784f);
    fma2c(p0, p1, p0, p1, z0, z1, 1.0f);
    mul2v(p0, p1, p0, p1, p0, p1);
    mul2v(p0, p1, p0, p1, p0, p1);
    mul2v(p0, p1, p0, p1, p0, p1);
    mul2v(p0, p1, p0, p1, p0, p1);
    float r0, r1;                                        // P^-16; P >= 1, and P^16 = inf gives 0: no range fix-ups needed
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r0) : "f"(p0));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r1) : "f"(p1));
    x0 = fmaf(-0.5f * a0, r0, fmaxf(x0, 0.f));
    x1 = fmaf(-0.5f * a1, r1, fmaxf(x1, 0.f));
}

// acc4 += p * x4 ;  acc4 *= a ;  dot of two float4 pairs
__device__ __forceinline__ void axpy4_2(float4& c, float p, const float4& x) { fma2(c.x, c.y, p, p, x.x, x.y); fma2(c.z, c.w, p, p, x.z, x.w); }
__device__ __forceinline__ void scale4_2(float4& c, float a) { mul2(c.x, c.y, a); mul2(c.z, c.w, a); }
__device__ __forceinline__ float dot8_2(const float4& qa, const float4& xa, const float4& qb, const float4& xb) {
    float sx = qa.x * xa.x, sy = qa.y * xa.y;
    fma2(sx, sy, qa.z, qa.w, xa.z, xa.w);
    fma2(sx, sy, qb.x, qb.y, xb.x, xb.y);
    fma2(sx, sy, qb.z, qb.w, xb.z, xb.w);
    return sx + sy;
}

__device__ __forceinline__ float half_warp_sum(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    return half_warp_sum(v);
}

// ---------------------------------------------------------------------------------------------
// mbarrier + 1-D bulk async copy (TMA engine, SASS UBLKCP) used to stream weight blobs
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// 1024-byte aligned start of the dynamic shared memory (what SWIZZLE_128B operands need), derived by pointer
// arithmetic from the array itself: an integer round trip would lose the shared address space and turn every
// later access into a generic LD / ST with 64-bit address arithmetic.
__device__ __forceinline__ unsigned char* smem_align1024(unsigned char* raw) {
    return raw + ((1024u - (smem_u32(raw) & 1023u)) & 1023u);
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
