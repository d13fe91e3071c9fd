// Weight gradients of the 64x64 Linear layers on the tensor cores (tcgen05), sm_100a.
//
// The reference gets dW = dY^T X from torch autograd inside MATTrainer.ppo_update (mat_trainer.py:146-147).
// Here the fused backward kernels (policy_bwd.cu) write each layer's output gradient dY [tok][64] to a scratch
// buffer; this kernel forms dW[out][in] = sum_tok dY[tok][out] * X[tok][in] for up to four layers per launch.
// K of the product is the token index.  The operands stay token-major: the tensor core reads them through
// MN-MAJOR shared-memory descriptors (a_major = b_major = 1), whose canonical SWIZZLE_128B layout -- 128-byte rows
// of 64 features per token, 8-token atoms of 1024 bytes, 16-byte chunk index XOR (token % 8) -- is exactly the
// layout the forward kernels write their activation operands in (tc::store_operand32): no transposition.
// fp16 hi + scaled-lo halves; the accumulators D1 = Ahi.Bhi, D2 = Ahi.Blo + Alo.Bhi stay in tensor memory over
// all chunks of a CTA and are added to the flat gradient with atomics at the end.  The MMA's M is 128: the second
// 64-feature atom of the A operand (reached through the descriptor's leading byte offset) is a block of zeros.
// X is rebuilt from what the forward saved: as stored, x_hat * ln_w + ln_b, or GELU(pre-activation).
// (scripts/tc/tc_test_dw.cu validated the accumulate-across-chunks scheme with transposed K-major operands.)
#include <stdlib.h>

#include "policy_bwd.h"
#include "tc.cuh"

using namespace tile;

namespace {

constexpr int TK = 64;                       // tokens per chunk (40 KB of operands: four CTAs per SM)
constexpr int OP_PART = TK * 128;            // bytes of one operand part: TK tokens x 64 features x 2 B
constexpr int A_BYTES = 3 * OP_PART;         // [hi][lo][zeros: the second 64-feature atom of both parts]
constexpr int B_BYTES = 2 * OP_PART;         // [hi][lo]
constexpr size_t DW_SMEM = 1024 + A_BYTES + B_BYTES + 2 * 64 * sizeof(float) + 64;

// MN-major SWIZZLE_128B descriptor: SBO = 1024 B between 8-token groups, LBO = bytes to the next 64-feature atom
__device__ __forceinline__ uint64_t make_desc_mn(uint32_t saddr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

__global__ void __launch_bounds__(NT, 4)
bwd_dw_kernel(const float* __restrict__ packed, float* __restrict__ grad, DwJobs jobs, long n_tok, int splits) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* base = smem_align1024(smem_raw);
    unsigned char* sA = base;                         // hi: [128 tokens][128 B], lo the same, then 16 KB of zeros
    unsigned char* sB = base + A_BYTES;               // hi: [128 tokens][128 B], then lo
    float* sLn = reinterpret_cast<float*>(sB + B_BYTES);         // LN weight[64] | bias[64] of the affine recipe
    uint64_t* bar = reinterpret_cast<uint64_t*>(sLn + 128);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int tid = threadIdx.x, warp = tid >> 5;
    const DwJob job = jobs.j[blockIdx.x / splits];
    const int split = blockIdx.x % splits;
    const long n_chunks = (n_tok + TK - 1) / TK;
    const long c0 = n_chunks * split / splits, c1 = n_chunks * (split + 1) / splits;
    if (c0 >= c1) return;
    if (tid == 0) { mbar_init(bar, 1); mbar_fence_init(); }
    if (warp == 0) tc::tmem_alloc_n<128>(tmem_slot);
    for (int i = tid; i < (A_BYTES + B_BYTES) / 16; i += NT) reinterpret_cast<uint4*>(base)[i] = make_uint4(0u, 0u, 0u, 0u);
    if (job.recipe == DW_X_AFFINE && tid < 128) sLn[tid] = packed[job.ln + lay::B_LNW + tid];
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *tmem_slot;
    const uint32_t idesc = (1u << 4) | (1u << 15) | (1u << 16) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);   // MN-major A and B
    const int hw = tid >> 4, l = tid & 15;            // half-warp hw loads token rows hw, hw + 16, ...; lane l: features 4l..4l+3
    float4 lw = make_float4(1.f, 1.f, 1.f, 1.f), lb = make_float4(0.f, 0.f, 0.f, 0.f);
    if (job.recipe == DW_X_AFFINE) { lw = *reinterpret_cast<const float4*>(sLn + l * 4); lb = *reinterpret_cast<const float4*>(sLn + 64 + l * 4); }
    uint32_t phase = 0;
    for (long c = c0; c < c1; ++c) {
        constexpr int NIT = TK / 16;
        float4 a4[NIT], x4[NIT];
#pragma unroll
        for (int it = 0; it < NIT; ++it) {             // coalesced: 256 bytes per half-warp and row
            const long g = c * TK + it * 16 + hw;
            const bool ok = g < n_tok;
            a4[it] = ok ? __ldg(reinterpret_cast<const float4*>(job.dY + g * 64) + l) : make_float4(0.f, 0.f, 0.f, 0.f);
            x4[it] = ok ? __ldg(reinterpret_cast<const float4*>(job.X + g * 64) + l) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int it = 0; it < NIT; ++it) {
            float4 x = x4[it];
            if (job.recipe == DW_X_AFFINE) {
                x.x = fmaf(x.x, lw.x, lb.x); x.y = fmaf(x.y, lw.y, lb.y); x.z = fmaf(x.z, lw.z, lb.z); x.w = fmaf(x.w, lw.w, lb.w);
            } else if (job.recipe == DW_X_GELU) {
                x.x = gelu_erf(x.x); x.y = gelu_erf(x.y); x.z = gelu_erf(x.z); x.w = gelu_erf(x.w);
            }
            tc::store_operand4(sA, it * 16 + hw, l, a4[it], OP_PART);   // hi at +0, lo at + OP_PART
            tc::store_operand4(sB, it * 16 + hw, l, x, OP_PART);        // rows past n_tok: dY = 0, so X does not matter
        }
        tc::publish();
        if (tid == 0) {
            const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
            const uint32_t first = c == c0 ? 0u : 1u;
#pragma unroll
            for (int ks = 0; ks < TK / 16; ++ks) {    // K-steps of 16 tokens = two 8-token atoms = 2048 bytes
                const uint32_t ko = ks * 2048;
                const uint32_t acc = (first || ks > 0) ? 1u : 0u;
                const uint64_t ah = make_desc_mn(a0 + ko, 2 * OP_PART - ko), al = make_desc_mn(a0 + OP_PART + ko, OP_PART - ko);   // LBO -> the zero block
                const uint64_t bh = make_desc_mn(b0 + ko, OP_PART), bl = make_desc_mn(b0 + OP_PART + ko, OP_PART);
                tc::mma_f16(tmem, ah, bh, idesc, acc);
                tc::mma_f16(tmem + 64, ah, bl, idesc, acc);
                tc::mma_f16(tmem + 64, al, bh, idesc, 1u);
            }
            tc::mma_commit(bar);
        }
        mbar_wait(bar, phase);                        // the next chunk overwrites the operands
        phase ^= 1;
        tc::fence_after_sync();
    }
    // rows m < 64 of the accumulator live in TMEM lanes 0..63: warps 0, 1 (columns 0-31) and 4, 5 (columns 32-63)
    if ((warp & 3) < 2) {
        const int m = (warp & 3) * 32 + (tid & 31), ch = warp >> 2;
        uint32_t v1[32], v2[32];
        const uint32_t taddr = tmem + ((uint32_t)((warp & 3) * 32) << 16) + ch * 32;
        TC_LD32(v1, taddr);
        TC_LD32(v2, taddr + 64);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        float* gp = grad + job.gW + m * 64 + ch * 32;
#pragma unroll
        for (int i = 0; i < 32; ++i) atomicAdd(gp + i, fmaf(__uint_as_float(v2[i]), 1.0f / 2048.0f, __uint_as_float(v1[i])));
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc_n<128>(tmem);
}

}  // namespace

int launch_bwd_dw(const float* packed, float* grad, const DwJobs& jobs, long n_tok, int n_sm, cudaStream_t st) {
    if (jobs.n <= 0) return MATAS_OK;
    CUDA_TRY(cudaFuncSetAttribute(bwd_dw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DW_SMEM));
    const long n_chunks = (n_tok + TK - 1) / TK;
    constexpr int per_sm = 3;         // CTAs per SM over all jobs of a launch; measured with up to six jobs per launch: 3 (5.43 ms backward) beats 2 (5.65) and 4 (5.53)
    long splits = ((long)per_sm * n_sm + jobs.n - 1) / jobs.n;
    if (splits > n_chunks) splits = n_chunks;
    if (splits < 1) splits = 1;
    bwd_dw_kernel<<<(unsigned)(jobs.n * splits), NT, DW_SMEM, st>>>(packed, grad, jobs, n_tok, (int)splits);
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}
