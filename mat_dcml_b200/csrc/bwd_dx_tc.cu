// Input gradient of the q / k / v projections on the tensor cores (tcgen05), sm_100a:
//     dX = dq.Wq + dk.Wk + dv.Wv          (KIND 0: q, k, v project the same input;  gX += dX)
//     dRep += dq.Wq ,  dX = dk.Wk + dv.Wv (KIND 1: decoder cross-attention, q from obs_rep;  gX = dX)
// plus the bias gradients (column sums of dq, dk, dv).  dq / dk / dv [tok][64] come from attn_bwd_tile.cu.
// A operands: the gradient tiles as token-major fp16 hi/lo halves, K-major (K = output feature), written with the
// same coalesced half-warp stores as everywhere else.  B operands: the FORWARD half blobs (rows = output feature,
// 128 bytes of input features each) read through an MN-major descriptor, i.e. B[n = in][k = out] = W[out][in]
// without a second weight image.  The three products of a tile accumulate in one TMEM accumulator.
// The epilogue stages the fp32 result in the (then idle) operand buffer so that the read-modify-write of the
// token-major gradient is coalesced.
#include "policy_bwd.h"
#include "tc.cuh"

using namespace tile;

namespace {

constexpr int WPART = 8192;                          // bytes of one weight operand part (64 rows x 128 B)
constexpr size_t DX_SMEM = 1024 + 2 * tc::A_PART + 3 * 2 * WPART + 3 * 16 * 64 * sizeof(float) + 64;

__device__ __forceinline__ uint64_t desc_mn(uint32_t saddr) {       // MN-major SWIZZLE_128B, one 64-element MN atom
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// fp32 staging inside the 32 KB operand buffer: row r (256 B), float4 chunk c stored at chunk c ^ (r & 15)
__device__ __forceinline__ float4* stage_at(unsigned char* op, int r, int c4) {
    return reinterpret_cast<float4*>(op + r * 256 + ((c4 ^ (r & 15)) << 4));
}

template <int KIND>
__global__ void __launch_bounds__(NT, 2)
bwd_qkv_dx_kernel(const float* __restrict__ packed, float* __restrict__ grad, DxArgs a, long n_tok) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* base = smem_align1024(smem_raw);
    unsigned char* op = base;                                     // [hi | lo][128 tokens][128 B]
    unsigned char* sW = base + 2 * tc::A_PART;                    // 3 x [hi | lo][64 out rows][128 B]
    float* sCol = reinterpret_cast<float*>(sW + 3 * 2 * WPART);   // [3][16 half-warps][64] bias-gradient partials
    uint64_t* bar = reinterpret_cast<uint64_t*>(sCol + 3 * 16 * 64);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int tid = threadIdx.x, warp = tid >> 5, hw = tid >> 4, l = tid & 15;
    if (tid == 0) { mbar_init(bar, 1); mbar_fence_init(); }
    if (warp == 0) tc::tmem_alloc_n<256>(tmem_slot);
    for (int m = 0; m < 3; ++m) {
        const uint4* src = reinterpret_cast<const uint4*>(packed + lay::hb(a.blob[m]));
        for (int i = tid; i < 2 * WPART / 16; i += NT) reinterpret_cast<uint4*>(sW + m * 2 * WPART)[i] = __ldg(src + i);
    }
    tc::fence_proxy_async();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *tmem_slot;
    // D f32, A f16 K-major, B f16 MN-major, N = 64, M = 128
    const uint32_t idesc = (1u << 4) | (1u << 16) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t a0 = smem_u32(op);
    const int row = (warp & 3) * 32 + (tid & 31), half = warp >> 2;          // epilogue ownership
    float4 cs[3];
    for (int m = 0; m < 3; ++m) cs[m] = make_float4(0.f, 0.f, 0.f, 0.f);
    uint32_t phase = 0;
    const long n_tiles = (n_tok + tc::TR - 1) / tc::TR;
    const float* src[3] = {a.dq, a.dk, a.dv};
    for (long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long tile0 = tile * tc::TR;
        const int n_valid = (int)min((long)tc::TR, n_tok - tile0);
#pragma unroll
        for (int m = 0; m < 3; ++m) {
            // gradient tile -> operand (coalesced), column sums for the bias gradient
            constexpr int NR = tc::TR * 16 / NT;          // rows per half-warp: all loads in flight before the first use
            float4 g4[NR];
#pragma unroll
            for (int i = 0; i < NR; ++i) {
                const int r = hw + i * (NT / 16);
                g4[i] = r < n_valid ? __ldg(reinterpret_cast<const float4*>(src[m] + (tile0 + r) * 64) + l) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int i = 0; i < NR; ++i) {
                cs[m].x += g4[i].x; cs[m].y += g4[i].y; cs[m].z += g4[i].z; cs[m].w += g4[i].w;
                tc::store_operand4(op, hw + i * (NT / 16), l, g4[i]);
            }
            tc::publish();
            if (tid == 0) {
                const uint32_t w0 = smem_u32(sW + m * 2 * WPART);
                // KIND 1: dq.Wq goes to its own accumulator (columns 128..255); dk.Wk starts the main one
                const uint32_t d = (KIND == 1 && m == 0) ? tmem + 128 : tmem;
                const bool fresh = m == 0 || (KIND == 1 && m == 1);
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {          // K = 16 output features per step: A + 32 B, B + 2 atoms of 8 rows
                    const uint32_t acc = (!fresh || ks > 0) ? 1u : 0u;
                    const uint64_t ah = tc::make_desc(a0 + ks * 32), al = tc::make_desc(a0 + tc::A_PART + ks * 32);
                    const uint64_t bh = desc_mn(w0 + ks * 2048), bl = desc_mn(w0 + WPART + ks * 2048);
                    tc::mma_f16(d, ah, bh, idesc, acc);
                    tc::mma_f16(d + 64, ah, bl, idesc, acc);
                    tc::mma_f16(d + 64, al, bh, idesc, 1u);
                }
                tc::mma_commit(bar);
            }
            mbar_wait(bar, phase);                       // the operand buffer is reused by the next product
            phase ^= 1;
            tc::fence_after_sync();
        }
        // ---- epilogue(s): TMEM -> fp32 staging in the operand buffer -> coalesced update of the gradient -----------
#pragma unroll
        for (int o = 0; o < (KIND == 1 ? 2 : 1); ++o) {
            uint32_t v1[32], v2[32];
            const uint32_t taddr = tmem + ((uint32_t)(row & ~31) << 16) + (o == 1 ? 128 : 0) + half * 32;
            TC_LD32(v1, taddr);
            TC_LD32(v2, taddr + 64);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float4 x;
                x.x = fmaf(__uint_as_float(v2[j * 4]), 1.0f / 2048.0f, __uint_as_float(v1[j * 4]));
                x.y = fmaf(__uint_as_float(v2[j * 4 + 1]), 1.0f / 2048.0f, __uint_as_float(v1[j * 4 + 1]));
                x.z = fmaf(__uint_as_float(v2[j * 4 + 2]), 1.0f / 2048.0f, __uint_as_float(v1[j * 4 + 2]));
                x.w = fmaf(__uint_as_float(v2[j * 4 + 3]), 1.0f / 2048.0f, __uint_as_float(v1[j * 4 + 3]));
                *stage_at(op, row, half * 8 + j) = x;
            }
            __syncthreads();
            float* dst = (o == 1) ? a.gRep : a.gX;
            const bool add = (o == 1) || KIND == 0;                 // KIND 1 main output overwrites gX
            constexpr int NR = tc::TR * 16 / NT;
            float4 g4[NR];                                          // the read-modify-write's reads, all in flight at once
#pragma unroll
            for (int i = 0; i < NR; ++i) {
                const int r = hw + i * (NT / 16);
                g4[i] = (add && r < n_valid) ? *(reinterpret_cast<const float4*>(dst + (tile0 + r) * 64) + l) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int i = 0; i < NR; ++i) {
                const int r = hw + i * (NT / 16);
                if (r < n_valid) {
                    const float4 x = *stage_at(op, r, l);
                    *(reinterpret_cast<float4*>(dst + (tile0 + r) * 64) + l) = make_float4(x.x + g4[i].x, x.y + g4[i].y, x.z + g4[i].z, x.w + g4[i].w);
                }
            }
            tc::fence_before_sync();
            __syncthreads();                                        // staging is free again (next epilogue / next tile's operand)
            tc::fence_after_sync();
        }
    }
    // ---- bias gradients: reduce the per-thread column sums over the 16 half-warps, one atomic per column ---------
    for (int m = 0; m < 3; ++m) *reinterpret_cast<float4*>(sCol + (m * 16 + hw) * 64 + l * 4) = cs[m];
    __syncthreads();
    if (tid < 192) {
        const int m = tid >> 6, c = tid & 63;
        float s = 0.f;
        for (int h = 0; h < 16; ++h) s += sCol[(m * 16 + h) * 64 + c];
        if (a.gb[m] >= 0) atomicAdd(grad + a.gb[m] + c, s);
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc_n<256>(tmem);
}

}  // namespace

int launch_bwd_qkv_dx(const float* packed, float* grad, const DxArgs& a, long n_tok, int kind, int n_sm, cudaStream_t st) {
    const long n_tiles = (n_tok + tc::TR - 1) / tc::TR;
    const unsigned grid = (unsigned)(n_tiles < 2L * n_sm ? n_tiles : 2L * n_sm);
    if (kind == 1) {
        CUDA_TRY(cudaFuncSetAttribute(bwd_qkv_dx_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DX_SMEM));
        bwd_qkv_dx_kernel<1><<<grid, NT, DX_SMEM, st>>>(packed, grad, a, n_tok);
    } else {
        CUDA_TRY(cudaFuncSetAttribute(bwd_qkv_dx_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DX_SMEM));
        bwd_qkv_dx_kernel<0><<<grid, NT, DX_SMEM, st>>>(packed, grad, a, n_tok);
    }
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}
