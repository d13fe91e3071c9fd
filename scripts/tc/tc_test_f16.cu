// Standalone check of the tcgen05 fp16 "hi + scaled lo" GEMM primitive:
//   x = hi + 2^-11 * lo, hi = fp16(x), lo = fp16((x - hi) * 2^11)
//   D1 += Ahi.Whi ; D2 += Ahi.Wlo + Alo.Whi ; D = D1 + 2^-11 * D2          (fp32 accumulators in TMEM)
// D[128x64] = A[128x64] . W[64x64]^T, fp32-class accuracy with 4 bytes of shared memory per operand element.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tc_test_f16.bin tc_test_f16.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t sw128_off(int row, int kbyte) {
    return (uint32_t)(row * 128 + ((((kbyte >> 4) ^ (row & 7)) << 4) | (kbyte & 15)));
}
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
__device__ __forceinline__ void mma_f16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n" ::"r"(tmem_d),
        "l"(da), "l"(db), "r"(idesc), "r"(accum), "r"(0u)
        : "memory");
}
__device__ __forceinline__ void split16(float x, __half& hi, __half& lo) {
    hi = __float2half_rn(x);
    lo = __float2half_rn((x - __half2float(hi)) * 2048.0f);
}
#define LD32(v, addr)                                                                                                          \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                                      \
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, " \
                 "%24, %25, %26, %27, %28, %29, %30, %31}, [%32];"                                                              \
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),  \
                   "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),       \
                   "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),      \
                   "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])                    \
                 : "r"(addr))

constexpr int M = 128, N = 64, K = 64;
constexpr int A_PART = M * 128, B_PART = N * 128;      // bytes: 64 halves = 128 B per row, one swizzle atom wide

__global__ void __launch_bounds__(256, 1) tc_gemm(const float* __restrict__ A, const float* __restrict__ W, float* __restrict__ D,
                                                  long long* cycles, int reps) {
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* sA = smem;                    // [hi|lo][128 rows][128 B]
    unsigned char* sB = smem + 2 * A_PART;       // [hi|lo][64 rows][128 B]
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" ::"r"(smem_u32(&tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (int idx = tid; idx < M * K; idx += 256) {
        const int m = idx / K, k = idx % K;
        __half hi, lo; split16(A[idx], hi, lo);
        const uint32_t off = sw128_off(m, k * 2);
        *reinterpret_cast<__half*>(sA + off) = hi;
        *reinterpret_cast<__half*>(sA + A_PART + off) = lo;
    }
    for (int idx = tid; idx < N * K; idx += 256) {
        const int n = idx / K, k = idx % K;
        __half hi, lo; split16(W[idx], hi, lo);
        const uint32_t off = sw128_off(n, k * 2);
        *reinterpret_cast<__half*>(sB + off) = hi;
        *reinterpret_cast<__half*>(sB + B_PART + off) = lo;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base;
    // instruction descriptor: D f32 (1<<4), A/B f16 (0), K-major, N = 64, M = 128
    const uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
    long long t0 = clock64();
    for (int rep = 0; rep < reps; ++rep) {
        if (tid == 0) {
            for (int ks = 0; ks < 4; ++ks)      // D1 = Ahi.Whi
                mma_f16(tmem_d, make_desc(a0 + ks * 32), make_desc(b0 + ks * 32), idesc, ks > 0);
            for (int ks = 0; ks < 4; ++ks)      // D2 = Ahi.Wlo
                mma_f16(tmem_d + 64, make_desc(a0 + ks * 32), make_desc(b0 + B_PART + ks * 32), idesc, ks > 0);
            for (int ks = 0; ks < 4; ++ks)      // D2 += Alo.Whi
                mma_f16(tmem_d + 64, make_desc(a0 + A_PART + ks * 32), make_desc(b0 + ks * 32), idesc, 1);
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        }
        asm volatile(
            "{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(
                smem_u32(&bar)),
            "r"(rep & 1)
            : "memory");
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
    long long t1 = clock64();
    if (tid == 0 && cycles) *cycles = t1 - t0;
    // epilogue: thread = (row = 32*(warp%4) + lane, column half = warp/4)
    uint32_t v1[32], v2[32];
    const int half = warp >> 2, row = (warp & 3) * 32 + (tid & 31);
    const uint32_t taddr = tmem_d + ((uint32_t)((warp & 3) * 32) << 16) + half * 32;
    LD32(v1, taddr);
    LD32(v2, taddr + 64);
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int c = 0; c < 32; ++c) D[row * 64 + half * 32 + c] = fmaf(__uint_as_float(v2[c]), 1.0f / 2048.0f, __uint_as_float(v1[c]));
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" ::"r"(tmem_d) : "memory");
}

int main() {
    float *hA = (float*)malloc(M * K * 4), *hW = (float*)malloc(N * K * 4), *hD = (float*)malloc(M * N * 4);
    float *dA, *dW, *dD; long long* dC;
    cudaMalloc(&dA, M * K * 4); cudaMalloc(&dW, N * K * 4); cudaMalloc(&dD, M * N * 4); cudaMalloc(&dC, 8);
    const int smem = 2 * A_PART + 2 * B_PART + 1024;
    cudaFuncSetAttribute(tc_gemm, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    int rc = 0;
    // scales: activations O(1)/O(10)/tiny, weights O(0.1)/O(1)
    const float sa[4] = {4.f, 40.f, 1e-3f, 4.f}, sw[4] = {1.f, 0.2f, 1.f, 1e-3f};
    for (int cse = 0; cse < 4; ++cse) {
        srand(1 + cse);
        for (int i = 0; i < M * K; ++i) hA[i] = (rand() / (float)RAND_MAX - 0.5f) * sa[cse];
        for (int i = 0; i < N * K; ++i) hW[i] = (rand() / (float)RAND_MAX - 0.5f) * sw[cse];
        cudaMemcpy(dA, hA, M * K * 4, cudaMemcpyHostToDevice); cudaMemcpy(dW, hW, N * K * 4, cudaMemcpyHostToDevice);
        cudaMemset(dD, 0, M * N * 4);
        tc_gemm<<<1, 256, smem>>>(dA, dW, dD, dC, 1);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("launch: %s\n", cudaGetErrorString(e)); return 1; }
        cudaMemcpy(hD, dD, M * N * 4, cudaMemcpyDeviceToHost);
        double max_err = 0, max_ref = 0, max_err32 = 0;
        for (int m = 0; m < M; ++m)
            for (int n = 0; n < N; ++n) {
                double ref = 0; float ref32 = 0.f;
                for (int k = 0; k < K; ++k) { ref += (double)hA[m * K + k] * hW[n * K + k]; ref32 = fmaf(hA[m * K + k], hW[n * K + k], ref32); }
                max_err = fmax(max_err, fabs(hD[m * N + n] - ref));
                max_err32 = fmax(max_err32, fabs((double)ref32 - ref));
                max_ref = fmax(max_ref, fabs(ref));
            }
        printf("case %d (|a|<=%g, |w|<=%g): tcgen05 f16 hi/lo: max |err| = %.3e (fp32 FMA chain: %.3e), max |ref| = %.4g\n", cse, sa[cse] / 2,
               sw[cse] / 2, max_err, max_err32, max_ref);
        if (!(max_err <= 4 * max_err32 + 1e-30)) rc = 2;
    }
    long long c1 = 0, c2 = 0;
    tc_gemm<<<1, 256, smem>>>(dA, dW, dD, dC, 1); cudaDeviceSynchronize(); cudaMemcpy(&c1, dC, 8, cudaMemcpyDeviceToHost);
    tc_gemm<<<1, 256, smem>>>(dA, dW, dD, dC, 101); cudaDeviceSynchronize(); cudaMemcpy(&c2, dC, 8, cudaMemcpyDeviceToHost);
    printf("issue->commit->wait round trip: %.0f cycles per 128x64x64 x3 GEMM (12 MMAs)\n", (c2 - c1) / 100.0);
    return rc;
}
