// Standalone check of the tcgen05 tf32x3 GEMM primitive: D[128x64] = A[128x64] . W[64x64]^T with fp32-class accuracy.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tc_test.bin tc_test.cu
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// K-major SWIZZLE_128B operand: rows of 128 bytes, 8-row atoms of 1024 bytes, 16-byte chunk index XOR (row % 8)
__device__ __forceinline__ uint32_t sw128_off(int row, int kbyte) {   // kbyte in [0,128)
    return (uint32_t)(row * 128 + ((((kbyte >> 4) ^ (row & 7)) << 4) | (kbyte & 15)));
}
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);            // start address
    d |= (uint64_t)1 << 16;                            // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;                  // stride byte offset: 8 rows * 128 B
    d |= (uint64_t)1 << 46;                            // descriptor version (Blackwell)
    d |= (uint64_t)2 << 61;                            // SWIZZLE_128B
    return d;
}
__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n" ::"r"(tmem_d),
        "l"(da), "l"(db), "r"(idesc), "r"(accum), "r"(0u)
        : "memory");
}
__device__ __forceinline__ float tf32_rna(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}

constexpr int M = 128, N = 64, K = 64;
constexpr int A_PANEL = M * 128, B_PANEL = N * 128;          // bytes per 32-element K panel
constexpr int A_PART = 2 * A_PANEL, B_PART = 2 * B_PANEL;    // hi or lo part (two panels)

__global__ void __launch_bounds__(128, 1) tc_gemm(const float* __restrict__ A, const float* __restrict__ W, float* __restrict__ D) {
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* sA = smem;                     // [hi|lo][panel][128 rows][128 B]
    unsigned char* sB = smem + 2 * A_PART;        // [hi|lo][panel][64 rows][128 B]
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(&tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // split + swizzled store
    for (int idx = tid; idx < M * K; idx += 128) {
        const int m = idx / K, k = idx % K;
        const float x = A[idx], hi = tf32_rna(x), lo = x - hi;
        const uint32_t off = (k / 32) * A_PANEL + sw128_off(m, (k % 32) * 4);
        *reinterpret_cast<float*>(sA + off) = hi;
        *reinterpret_cast<float*>(sA + A_PART + off) = lo;
    }
    for (int idx = tid; idx < N * K; idx += 128) {
        const int n = idx / K, k = idx % K;
        const float x = W[idx], hi = tf32_rna(x), lo = x - hi;
        const uint32_t off = (k / 32) * B_PANEL + sw128_off(n, (k % 32) * 4);
        *reinterpret_cast<float*>(sB + off) = hi;
        *reinterpret_cast<float*>(sB + B_PART + off) = lo;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // generic-proxy writes -> async proxy (MMA) reads
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base;
    if (tid == 0) {
        // instruction descriptor: D f32, A/B tf32, both K-major, N = 64, M = 128
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
        uint32_t accum = 0;
        // small cross terms first, then hi*hi
        const int pa[3] = {1, 0, 0}, pb[3] = {0, 1, 0};       // part index: 0 = hi, 1 = lo
        for (int t = 0; t < 3; ++t)
            for (int p = 0; p < 2; ++p)
                for (int ks = 0; ks < 4; ++ks) {
                    const uint64_t da = make_desc(a0 + pa[t] * A_PART + p * A_PANEL + ks * 32);
                    const uint64_t db = make_desc(b0 + pb[t] * B_PART + p * B_PANEL + ks * 32);
                    mma_tf32(tmem_d, da, db, idesc, accum);
                    accum = 1;
                }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    // wait for the MMAs
    asm volatile(
        "{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(
            smem_u32(&bar)),
        "r"(0)
        : "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // epilogue: thread = row; 64 columns
    uint32_t v[64];
    const uint32_t taddr = tmem_d + ((uint32_t)(warp * 32) << 16);
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, "
        "%29, %30, %31, %32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, %48, %49, %50, %51, %52, %53, %54, %55, "
        "%56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
          "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
          "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]),
          "=r"(v[31]), "=r"(v[32]), "=r"(v[33]), "=r"(v[34]), "=r"(v[35]), "=r"(v[36]), "=r"(v[37]), "=r"(v[38]), "=r"(v[39]), "=r"(v[40]),
          "=r"(v[41]), "=r"(v[42]), "=r"(v[43]), "=r"(v[44]), "=r"(v[45]), "=r"(v[46]), "=r"(v[47]), "=r"(v[48]), "=r"(v[49]), "=r"(v[50]),
          "=r"(v[51]), "=r"(v[52]), "=r"(v[53]), "=r"(v[54]), "=r"(v[55]), "=r"(v[56]), "=r"(v[57]), "=r"(v[58]), "=r"(v[59]), "=r"(v[60]),
          "=r"(v[61]), "=r"(v[62]), "=r"(v[63])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    const int row = tid;
    for (int c = 0; c < 64; ++c) D[row * 64 + c] = __uint_as_float(v[c]);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem_d) : "memory");
}

int main() {
    float *hA = (float*)malloc(M * K * 4), *hW = (float*)malloc(N * K * 4), *hD = (float*)malloc(M * N * 4);
    srand(1);
    for (int i = 0; i < M * K; ++i) hA[i] = (rand() / (float)RAND_MAX - 0.5f) * 4.f;
    for (int i = 0; i < N * K; ++i) hW[i] = (rand() / (float)RAND_MAX - 0.5f) * 1.f;
    float *dA, *dW, *dD;
    cudaMalloc(&dA, M * K * 4); cudaMalloc(&dW, N * K * 4); cudaMalloc(&dD, M * N * 4);
    cudaMemcpy(dA, hA, M * K * 4, cudaMemcpyHostToDevice); cudaMemcpy(dW, hW, N * K * 4, cudaMemcpyHostToDevice);
    cudaMemset(dD, 0, M * N * 4);
    const int smem = 2 * A_PART + 2 * B_PART + 1024;
    cudaFuncSetAttribute(tc_gemm, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    tc_gemm<<<1, 128, smem>>>(dA, dW, dD);
    cudaError_t e = cudaDeviceSynchronize();
    printf("launch: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
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
    printf("tcgen05 tf32x3: max |err| = %.3e (fp32 FMA chain: %.3e), max |ref| = %.3f, D[0..3] = %f %f %f %f\n", max_err, max_err32, max_ref,
           hD[0], hD[1], hD[2], hD[3]);
    return max_err < 1e-4 ? 0 : 2;
}
