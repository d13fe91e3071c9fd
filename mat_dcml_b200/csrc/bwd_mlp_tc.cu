// Backward of a block tail with its products on the tensor cores (tcgen05), sm_100a:
//   [ LayerNorm_F backward -> mlp.2 -> GELU backward -> mlp.0 -> ] LayerNorm_M backward -> residual gradient ->
//   projection backward (dY of the attention output, D = rowsum_head(dY * Y))
// i.e. the work of bwd_mlp_proj (policy_bwd.cu) for the encoder block tail and the two decoder block halves
// (autograd of EncodeBlock / DecodeBlock, ma_transformer.py:70-116).  The element-wise and LayerNorm parts stay in
// the coalesced half-warp-per-row layout (lane l: columns 4l..4l+3; row sums by shuffles; parameter / bias gradients
// as per-thread column sums).  Each dY.W product takes the gradient tile as a K-major fp16 hi/lo operand and the
// FORWARD half blob of the layer as an MN-major B operand; the result is read from tensor memory thread-per-row
// and brought back to the half-warp layout through a swizzled fp32 staging pass in the operand buffer.
// The layer-output gradients (dS, dM0, d proj) are also written to a.sY for the weight-gradient kernel.
#include "policy_bwd.h"
#include "tc.cuh"

using namespace tile;

namespace {

constexpr int WPART = 8192;
constexpr int NCOL = 7;                              // column-sum accumulators per thread
constexpr size_t MLP_SMEM = 1024 + 2 * tc::A_PART + 3 * 2 * WPART + 64;      // the column-sum reduction reuses the operand buffer
static_assert(NCOL * 16 * 64 * sizeof(float) <= 2 * tc::A_PART, "column-sum scratch must fit the operand buffer");

__device__ __forceinline__ uint64_t desc_mn(uint32_t saddr) {       // MN-major SWIZZLE_128B, one 64-element MN atom
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
__device__ __forceinline__ float4* stage_at(unsigned char* op, int r, int c4) {
    return reinterpret_cast<float4*>(op + r * 256 + ((c4 ^ (r & 15)) << 4));
}
__device__ __forceinline__ float gelu_grad(float h) {
    const float cdf = 0.5f * (1.0f + erf_as(h * 0.70710678118654752f));       // A&S 7.1.26: 1.5e-7 absolute
    return cdf + h * 0.3989422804014327f * expf(-0.5f * h * h);
}
__device__ __forceinline__ void add4(float4& a, const float4& b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
__device__ __forceinline__ float4 mul4(const float4& a, const float4& b) { return make_float4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
__device__ __forceinline__ float4 ld4g(const float* p, long row, int l) { return __ldg(reinterpret_cast<const float4*>(p + row * 64) + l); }

// LayerNorm backward of one row chunk: d -> dx; accumulates d(weight) += d * xh and d(bias) += d
__device__ __forceinline__ float4 ln_bwd4(const float4& d, const float4& xh, float rs, const float4& w, float4& gw, float4& gb) {
    add4(gw, mul4(d, xh));
    add4(gb, d);
    const float4 g = mul4(d, w);
    const float s1 = half_warp_sum(g.x + g.y + g.z + g.w) * (1.0f / 64.0f);
    const float s2 = half_warp_sum(g.x * xh.x + g.y * xh.y + g.z * xh.z + g.w * xh.w) * (1.0f / 64.0f);
    return make_float4(rs * (g.x - s1 - xh.x * s2), rs * (g.y - s1 - xh.y * s2), rs * (g.z - s1 - xh.z * s2), rs * (g.w - s1 - xh.w * s2));
}

struct Mma {
    uint32_t tmem, a0, idesc;
    uint64_t* bar;
    uint32_t phase;
    // publish the operand, one thread issues D = A . W (W: forward half blob at shared address w), all wait
    __device__ __forceinline__ void run(uint32_t w) {
        tc::publish();
        if (threadIdx.x == 0) {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                const uint32_t acc = ks > 0 ? 1u : 0u;
                const uint64_t ah = tc::make_desc(a0 + ks * 32), al = tc::make_desc(a0 + tc::A_PART + ks * 32);
                const uint64_t bh = desc_mn(w + ks * 2048), bl = desc_mn(w + WPART + ks * 2048);
                tc::mma_f16(tmem, ah, bh, idesc, acc);
                tc::mma_f16(tmem + 64, ah, bl, idesc, acc);
                tc::mma_f16(tmem + 64, al, bh, idesc, 1u);
            }
            tc::mma_commit(bar);
        }
        mbar_wait(bar, phase);
        phase ^= 1;
        tc::fence_after_sync();
    }
    // TMEM -> fp32 staging (swizzled rows of 256 B in the operand buffer); ends with a CTA barrier
    __device__ __forceinline__ void stage(unsigned char* op) {
        const int warp = threadIdx.x >> 5, row = (warp & 3) * 32 + (threadIdx.x & 31), half = warp >> 2;
        uint32_t v1[32], v2[32];
        const uint32_t taddr = tmem + ((uint32_t)(row & ~31) << 16) + half * 32;
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
        tc::fence_before_sync();
        __syncthreads();
        tc::fence_after_sync();
    }
};

template <bool WITH_MLP>
__global__ void __launch_bounds__(NT, 2)
bwd_mlp_tc_kernel(const float* __restrict__ packed, float* __restrict__ grad, TailArgs a, long n_tok) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* base = smem_align1024(smem_raw);
    unsigned char* op = base;                                     // operand [hi | lo][128][128 B]  /  fp32 staging [128][256 B]
    unsigned char* sW = base + 2 * tc::A_PART;                    // forward half blobs of mlp.2, mlp.0, proj
    uint64_t* bar = reinterpret_cast<uint64_t*>(sW + 3 * 2 * WPART);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int tid = threadIdx.x, warp = tid >> 5, hw = tid >> 4, l = tid & 15;
    if (tid == 0) { mbar_init(bar, 1); mbar_fence_init(); }
    if (warp == 0) tc::tmem_alloc_n<128>(tmem_slot);
    const int blobs[3] = {a.lnF, a.fW0, a.lnM};                   // mlp.2 lives in the lnF blob, proj in the lnM blob
    for (int m = WITH_MLP ? 0 : 2; m < 3; ++m) {
        const uint4* src = reinterpret_cast<const uint4*>(packed + lay::hb(blobs[m]));
        for (int i = tid; i < 2 * WPART / 16; i += NT) reinterpret_cast<uint4*>(sW + m * 2 * WPART)[i] = __ldg(src + i);
    }
    tc::fence_proxy_async();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    Mma mma;
    mma.tmem = *tmem_slot; mma.a0 = smem_u32(op); mma.bar = bar; mma.phase = 0;
    mma.idesc = (1u << 4) | (1u << 16) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);   // A K-major, B MN-major
    const uint32_t w2 = smem_u32(sW), w0 = smem_u32(sW + 2 * WPART), wp = smem_u32(sW + 4 * WPART);
    const float4 lnFw = WITH_MLP ? __ldg(reinterpret_cast<const float4*>(packed + a.lnF + lay::B_LNW) + l) : make_float4(0.f, 0.f, 0.f, 0.f);
    const float4 lnMw = __ldg(reinterpret_cast<const float4*>(packed + a.lnM + lay::B_LNW) + l);
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 cs[NCOL];                                              // gFw gFb db2 db0 gMw gMb dbp
#pragma unroll
    for (int i = 0; i < NCOL; ++i) cs[i] = z4;
    const long n_tiles = (n_tok + tc::TR - 1) / tc::TR;
    for (long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long tile0 = tile * tc::TR;
        float4 t[8];
        if (WITH_MLP) {
            // ---- LayerNorm_F backward -> dS ------------------------------------------------------------------
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int r = hw + 16 * i;
                const long g = tile0 + r;
                const bool ok = g < n_tok;
                const float4 d = ok ? ld4g(a.gOut, g, l) : z4, xh = ok ? ld4g(a.xhF, g, l) : z4;
                const float rs = ok ? __ldg(a.rsF + g) : 0.f;
                const float4 ds = ln_bwd4(d, xh, rs, lnFw, cs[0], cs[1]);
                add4(cs[2], ds);
                if (ok) *(reinterpret_cast<float4*>(a.sY[0] + g * 64) + l) = ds;
                tc::store_operand4(op, r, l, ds);
            }
            mma.run(w2);                                          // dG = dS . W2
            mma.stage(op);
#pragma unroll
            for (int i = 0; i < 8; ++i) t[i] = *stage_at(op, hw + 16 * i, l);
            __syncthreads();
            // ---- GELU backward -> dM0 ------------------------------------------------------------------------
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int r = hw + 16 * i;
                const long g = tile0 + r;
                const bool ok = g < n_tok;
                const float4 h = ok ? ld4g(a.m0, g, l) : z4;
                const float4 dm = make_float4(t[i].x * gelu_grad(h.x), t[i].y * gelu_grad(h.y), t[i].z * gelu_grad(h.z), t[i].w * gelu_grad(h.w));
                const float4 dmz = ok ? dm : z4;
                add4(cs[3], dmz);
                if (ok) *(reinterpret_cast<float4*>(a.sY[1] + g * 64) + l) = dmz;
                tc::store_operand4(op, r, l, dmz);
            }
            mma.run(w0);                                          // dM0 . W0
            mma.stage(op);
#pragma unroll
            for (int i = 0; i < 8; ++i) t[i] = *stage_at(op, hw + 16 * i, l);
            __syncthreads();
        }
        // ---- LayerNorm_M backward -> gradient of (residual + proj) ---------------------------------------------
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int r = hw + 16 * i;
            const long g = tile0 + r;
            const bool ok = g < n_tok;
            float4 d;
            if (WITH_MLP) {                                       // dX_M = dS + dM0 . W0
                d = ok ? *(reinterpret_cast<const float4*>(a.sY[0] + g * 64) + l) : z4;
                add4(d, t[i]);
                if (!ok) d = z4;
            } else {
                d = ok ? ld4g(a.gOut, g, l) : z4;
            }
            const float4 xh = ok ? ld4g(a.xhM, g, l) : z4;
            const float rs = ok ? __ldg(a.rsM + g) : 0.f;
            const float4 dp = ln_bwd4(d, xh, rs, lnMw, cs[4], cs[5]);
            add4(cs[6], dp);
            if (ok) {
                float4* gp = reinterpret_cast<float4*>(a.gRes + g * 64) + l;
                float4 o = dp;
                if (a.res_add) add4(o, *gp);
                *gp = o;
                *(reinterpret_cast<float4*>(a.sY[2] + g * 64) + l) = dp;
            }
            tc::store_operand4(op, r, l, dp);
        }
        mma.run(wp);                                              // dY = d . Wp
        mma.stage(op);
#pragma unroll
        for (int i = 0; i < 8; ++i) t[i] = *stage_at(op, hw + 16 * i, l);
        __syncthreads();
        // ---- attention-output gradient and D = rowsum_head(dY * Y) ------------------------------------------------
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const long g = tile0 + hw + 16 * i;
            const bool ok = g < n_tok;
            const float4 y = ok ? ld4g(a.Y, g, l) : z4;
            float p = t[i].x * y.x + t[i].y * y.y + t[i].z * y.z + t[i].w * y.w;
            p += __shfl_xor_sync(0xffffffffu, p, 4);
            p += __shfl_xor_sync(0xffffffffu, p, 2);
            p += __shfl_xor_sync(0xffffffffu, p, 1);
            if (ok) {
                *(reinterpret_cast<float4*>(a.gY + g * 64) + l) = t[i];
                if ((l & 7) == 0) a.gD[g * 2 + (l >> 3)] = p;
            }
        }
    }
    // ---- parameter / bias gradients: reduce the per-thread column sums over the 16 half-warps --------------------
    __syncthreads();
    float* sCol = reinterpret_cast<float*>(op);                   // [NCOL][16][64]
#pragma unroll
    for (int m = 0; m < NCOL; ++m) *reinterpret_cast<float4*>(sCol + (m * 16 + hw) * 64 + l * 4) = cs[m];
    __syncthreads();
    const int offs[NCOL] = {a.glnFw, a.glnFb, a.gb2, a.gb0, a.glnMw, a.glnMb, a.gbp};
    for (int idx = tid; idx < NCOL * 64; idx += NT) {
        const int m = idx >> 6, c = idx & 63;
        if (!WITH_MLP && m < 4) continue;
        float sum = 0.f;
        for (int h = 0; h < 16; ++h) sum += sCol[(m * 16 + h) * 64 + c];
        if (offs[m] >= 0) atomicAdd(grad + offs[m] + c, sum);
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc_n<128>(mma.tmem);
}

}  // namespace

int launch_bwd_mlp_tc(const float* packed, float* grad, const TailArgs& a, long n_tok, int n_sm, cudaStream_t st) {
    const long n_tiles = (n_tok + tc::TR - 1) / tc::TR;
    const unsigned grid = (unsigned)(n_tiles < 2L * n_sm ? n_tiles : 2L * n_sm);
    if (a.with_mlp) {
        CUDA_TRY(cudaFuncSetAttribute(bwd_mlp_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MLP_SMEM));
        bwd_mlp_tc_kernel<true><<<grid, NT, MLP_SMEM, st>>>(packed, grad, a, n_tok);
    } else {
        CUDA_TRY(cudaFuncSetAttribute(bwd_mlp_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MLP_SMEM));
        bwd_mlp_tc_kernel<false><<<grid, NT, MLP_SMEM, st>>>(packed, grad, a, n_tok);
    }
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}
