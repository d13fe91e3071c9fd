// Standalone check of the weight-gradient product on tcgen05:  D[m][n] = sum_t dY[t][m] * X1[t][n],
// m < 64 (padded to the MMA's M = 128 with zero rows), X1 = [X | 1] (n < 72: 64 inputs, a ones column for the bias
// gradient, 7 zero columns), K = tokens.  Operands are written TRANSPOSED (token index = K, contiguous) as fp16
// hi + scaled-lo halves in the K-major SWIZZLE_128B layout, two 64-token panels per 128-token chunk; the accumulator
// stays in tensor memory across chunks.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tc_test_dw.bin tc_test_dw.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
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
#define LD8(v, addr)                                                                                   \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"       \
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(addr))

constexpr int TK = 128;                    // tokens per chunk
constexpr int NB = 72;                     // B rows: 64 inputs + ones + 7 zero
constexpr int A_PANEL = 128 * 128;         // bytes: 128 rows (64 real) x 64 tokens x 2 B
constexpr int A_PART = 2 * A_PANEL;
constexpr int B_PANEL = NB * 128;
constexpr int B_PART = 2 * B_PANEL;
constexpr int DCOLS = 256;                 // TMEM: D1 at 0..71, D2 at 128..199

// element (row r = feature, token t of the chunk): byte offset inside one part
__device__ __forceinline__ uint32_t t_off(int r, int t, int panel_bytes) {
    const int kk = t & 63;
    return (uint32_t)((t >> 6) * panel_bytes + r * 128 + ((((kk >> 3) ^ (r & 7)) << 4) | ((kk & 7) * 2)));
}

__global__ void __launch_bounds__(256, 1) dw_kernel(const float* __restrict__ dY, const float* __restrict__ X, float* __restrict__ D,
                                                    int n_chunks) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* base = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    unsigned char* sA = base;                    // [hi | lo][panel 2][128 rows][128 B]
    unsigned char* sB = base + 2 * A_PART;       // [hi | lo][panel 2][72 rows][128 B]   (72 * 128 = 9216 = 9 * 1024)
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(&tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (int i = tid; i < (2 * A_PART + 2 * B_PART) / 16; i += 256) reinterpret_cast<uint4*>(base)[i] = make_uint4(0u, 0u, 0u, 0u);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base;
    const uint32_t idesc = (1u << 4) | ((uint32_t)(NB >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const int t = tid & 127, half = tid >> 7;                // this thread: token t of the chunk, features 32*half ..
    // the ones row of B (row 64) never changes
    if (half == 0) {
        const __half one = __float2half_rn(1.0f);
        *reinterpret_cast<__half*>(sB + t_off(64, t, B_PANEL)) = one;
    }
    uint32_t phase = 0;
    for (int c = 0; c < n_chunks; ++c) {
        const float* dy = dY + ((long)c * TK + t) * 64 + half * 32;
        const float* x = X + ((long)c * TK + t) * 64 + half * 32;
        for (int j = 0; j < 32; ++j) {
            const int r = half * 32 + j;
            const float a = dy[j], b = x[j];
            const __half ah = __float2half_rn(a), bh = __float2half_rn(b);
            const __half al = __float2half_rn((a - __half2float(ah)) * 2048.0f), bl = __float2half_rn((b - __half2float(bh)) * 2048.0f);
            const uint32_t oa = t_off(r, t, A_PANEL), ob = t_off(r, t, B_PANEL);
            *reinterpret_cast<__half*>(sA + oa) = ah;
            *reinterpret_cast<__half*>(sA + A_PART + oa) = al;
            *reinterpret_cast<__half*>(sB + ob) = bh;
            *reinterpret_cast<__half*>(sB + B_PART + ob) = bl;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (tid == 0) {
            const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
            for (int ks = 0; ks < 8; ++ks) {           // 8 K-steps of 16 tokens: panel ks / 4, 32-byte step inside
                const uint32_t ao = (ks >> 2) * A_PANEL + (ks & 3) * 32, bo = (ks >> 2) * B_PANEL + (ks & 3) * 32;
                mma_f16(tmem, make_desc(a0 + ao), make_desc(b0 + bo), idesc, (c > 0 || ks > 0) ? 1u : 0u);                    // D1 += Ahi.Bhi
                mma_f16(tmem + 128, make_desc(a0 + ao), make_desc(b0 + B_PART + bo), idesc, (c > 0 || ks > 0) ? 1u : 0u);    // D2 += Ahi.Blo
                mma_f16(tmem + 128, make_desc(a0 + A_PART + ao), make_desc(b0 + bo), idesc, 1u);                             // D2 += Alo.Bhi
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        }
        // the operands are overwritten by the next chunk: wait for the MMAs
        asm volatile(
            "{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(
                smem_u32(&bar)),
            "r"(phase)
            : "memory");
        phase ^= 1;
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
    // epilogue: rows m < 64 live in TMEM lanes 0..63 -> warps 0, 1 (and 4, 5 for the second column half)
    if ((warp & 3) < 2) {
        const int m = (warp & 3) * 32 + (tid & 31), chalf = warp >> 2;      // columns [40*chalf, 40*chalf + 40) of 72 -> use 8-wide loads
        for (int cb = chalf * 40; cb < chalf * 40 + 40 && cb < NB; cb += 8) {
            uint32_t v1[8], v2[8];
            const uint32_t taddr = tmem + ((uint32_t)((warp & 3) * 32) << 16) + cb;
            LD8(v1, taddr);
            LD8(v2, taddr + 128);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            for (int i = 0; i < 8; ++i)
                if (cb + i < 65) D[m * 65 + cb + i] = fmaf(__uint_as_float(v2[i]), 1.0f / 2048.0f, __uint_as_float(v1[i]));
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem) : "memory");
}

int main() {
    const int n_chunks = 5, T = n_chunks * TK;
    float *hY = (float*)malloc(T * 64 * 4), *hX = (float*)malloc(T * 64 * 4), *hD = (float*)malloc(64 * 65 * 4);
    srand(3);
    for (int i = 0; i < T * 64; ++i) { hY[i] = (rand() / (float)RAND_MAX - 0.5f) * 0.02f; hX[i] = (rand() / (float)RAND_MAX - 0.5f) * 4.f; }
    float *dY, *dX, *dD;
    cudaMalloc(&dY, T * 64 * 4); cudaMalloc(&dX, T * 64 * 4); cudaMalloc(&dD, 64 * 65 * 4);
    cudaMemcpy(dY, hY, T * 64 * 4, cudaMemcpyHostToDevice); cudaMemcpy(dX, hX, T * 64 * 4, cudaMemcpyHostToDevice);
    cudaMemset(dD, 0, 64 * 65 * 4);
    const int smem = 2 * A_PART + 2 * B_PART + 1024;
    cudaFuncSetAttribute(dw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    dw_kernel<<<1, 256, smem>>>(dY, dX, dD, n_chunks);
    cudaError_t e = cudaDeviceSynchronize();
    printf("launch: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    cudaMemcpy(hD, dD, 64 * 65 * 4, cudaMemcpyDeviceToHost);
    double max_err = 0, max_ref = 0, max_err32 = 0;
    for (int m = 0; m < 64; ++m)
        for (int n = 0; n < 65; ++n) {
            double ref = 0; float ref32 = 0.f;
            for (int t = 0; t < T; ++t) {
                const float xv = n < 64 ? hX[t * 64 + n] : 1.0f;
                ref += (double)hY[t * 64 + m] * xv; ref32 = fmaf(hY[t * 64 + m], xv, ref32);
            }
            max_err = fmax(max_err, fabs(hD[m * 65 + n] - ref));
            max_err32 = fmax(max_err32, fabs((double)ref32 - ref));
            max_ref = fmax(max_ref, fabs(ref));
        }
    printf("tcgen05 dW = dY^T [X|1] over %d tokens: max |err| = %.3e (fp32 FMA chain: %.3e), max |ref| = %.4g, D[0][0..2] = %g %g %g, db[0] = %g\n", T,
           max_err, max_err32, max_ref, hD[0], hD[1], hD[2], hD[64]);
    return max_err <= 4 * max_err32 + 1e-9 ? 0 : 2;
}
