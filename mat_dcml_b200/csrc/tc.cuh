// tcgen05 building blocks for the [128 tokens x 64] . [64 x 64] stages of the token-parallel kernels (sm_100a).
//
// Precision scheme ("fp16 hi + scaled lo", validated standalone in scripts/tc/tc_test_f16.cu: error below the
// fp32 FMA chain's): every operand element x is stored as two halves, hi = fp16(x), lo = fp16((x - hi) * 2^11),
// and a GEMM is three fp16 MMAs with fp32 accumulation in tensor memory,
//      D1 = Ahi.Whi          D2 = Ahi.Wlo + Alo.Whi          D = D1 + 2^-11 D2
// (the dropped lo.lo term is 2^-22 relative).  Activations on this path are LayerNorm / GELU / softmax-average
// outputs and the weights are O(1), far inside fp16's range; values beyond 65504 would overflow.
//
// Shared-memory operands are K-major with the 128-byte swizzle: row r (token or output feature) occupies
// 128 bytes = 64 halves at r*128, its 16-byte chunk c stored at chunk position c ^ (r % 8).
// Thread layout of the epilogues: 256 threads, thread t owns row 32*(warp % 4) + lane (the TMEM lane its warp
// may read) and the column half warp / 4 (32 accumulator columns), so LayerNorm needs one exchange with the
// partner thread through shared memory and no shuffles.
#pragma once
#include <cuda_fp16.h>

#include "tile.cuh"

namespace tc {
using namespace tile;

constexpr int TR = 128;                       // rows (tokens) per tile
constexpr int A_PART = TR * 128;              // bytes of one operand part (hi or lo) of the activation tile
constexpr int SLOT_BYTES = 17408;             // ring slot stride: 1024-byte aligned, holds one 17152-byte half blob
constexpr int SLOT_F = SLOT_BYTES / 4;
constexpr int TILE_BYTES = TR * XS * 4;       // fp32 row-major staging / residual tile
constexpr int TMEM_COLS = 128;                // D1: columns 0-63, D2: columns 64-127
constexpr int MMA_BAR = 7;                    // index of the MMA-completion barrier in Smem::bars

struct Smem {
    unsigned char* sA;     // [hi | lo][128 rows][128 B]
    float* ring;           // depth slots of SLOT_F floats
    float* sT;             // [128][XS] fp32
    float* sSmall;         // 1024 floats
    uint64_t* bars;        // [0, depth): weight ring, [MMA_BAR]: MMA completion
    uint32_t* tmem_slot;
};
__host__ __device__ constexpr size_t smem_bytes(int depth) {
    return 1024 + 2 * (size_t)A_PART + (size_t)depth * SLOT_BYTES + TILE_BYTES + 4096 + 8 * 8 + 16;
}
__device__ inline Smem carve(unsigned char* raw, int depth) {
    Smem s;
    unsigned char* p = smem_align1024(raw);
    s.sA = p; p += 2 * A_PART;
    s.ring = reinterpret_cast<float*>(p); p += (size_t)depth * SLOT_BYTES;
    s.sT = reinterpret_cast<float*>(p); p += TILE_BYTES;
    s.sSmall = reinterpret_cast<float*>(p); p += 4096;
    s.bars = reinterpret_cast<uint64_t*>(p); p += 8 * 8;
    s.tmem_slot = reinterpret_cast<uint32_t*>(p);
    return s;
}

// ---- raw tcgen05 / proxy helpers ------------------------------------------------------------------
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t* slot) {   // one full warp
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "n"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t tmem) {  // one full warp
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TMEM_COLS) : "memory");
}
// K-major SWIZZLE_128B shared-memory matrix descriptor: 8-row atoms of 1024 bytes
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// Descriptors of the K = 16 slices of one operand differ only in the start-address field: build the low word once
// (desc_lo) and step it by 32 bytes >> 4 = 2 per slice; the high word is a constant.  Operands are 1024-byte aligned,
// so the step never carries out of the 14-bit field.
__device__ __forceinline__ uint32_t desc_lo(uint32_t saddr) { return ((saddr >> 4) & 0x3fffu) | (1u << 16); }
constexpr uint32_t DESC_HI = (uint32_t)(1024 >> 4) | (1u << 14) | (2u << 29);
__device__ __forceinline__ uint64_t desc_at(uint32_t lo, int ks) { return ((uint64_t)DESC_HI << 32) | (uint64_t)(lo + 2u * (uint32_t)ks); }
__device__ __forceinline__ void mma_f16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n" ::"r"(tmem_d),
        "l"(da), "l"(db), "r"(idesc), "r"(accum), "r"(0u)
        : "memory");
}
#define TC_LD32(v, addr)                                                                                                       \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                                      \
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, " \
                 "%24, %25, %26, %27, %28, %29, %30, %31}, [%32];"                                                              \
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),  \
                   "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),       \
                   "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),      \
                   "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])                    \
                 : "r"(addr))

#define TC_LD16(v, addr)                                                                                                       \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 "                                                                      \
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"                                \
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),  \
                   "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])                     \
                 : "r"(addr))

#define TC_ST16(addr, v)                                                                                                       \
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "                                                                \
                 "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"                                      \
                 ::"r"(addr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),  \
                   "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])                            \
                 : "memory")

template <int COLS>
__device__ __forceinline__ void tmem_alloc_n(uint32_t* slot) {   // one full warp
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "n"(COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int COLS>
__device__ __forceinline__ void tmem_dealloc_n(uint32_t tmem) {  // one full warp
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(COLS) : "memory");
}
// one elected thread: (D1, D2) at TMEM columns [d, d+64), [d+64, d+128) (+)= A . W for the fp16 hi/lo operand at
// shared address a (lo part at + A_PART) and the half blob at shared address w
__device__ __forceinline__ void mma_set(uint32_t d, uint32_t a, uint32_t w, bool accumulate) {
    // The half blob holds Whi (64 rows) and Wlo (64 rows) back to back, i.e. ONE 128-row K-major tile, and D1 | D2 are adjacent
    // accumulator columns: Ahi . [Whi | Wlo] is a single N = 128 product (the A tile is read from shared memory once for
    // both), then Alo . Whi accumulates into D2: 8 MMA instructions per [128 x 64].[64 x 64] stage instead of 12.
    const uint32_t idesc128 = (1u << 4) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t idesc64 = (1u << 4) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    static_assert(lay::HB_LO_BYTES == 64 * 128, "Wlo must follow Whi as rows 64..127 of one swizzled tile");
    const uint32_t da = desc_lo(a), dl = desc_lo(a + A_PART), dw = desc_lo(w);
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) mma_f16(d, desc_at(da, ks), desc_at(dw, ks), idesc128, (accumulate || ks > 0) ? 1u : 0u);
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) mma_f16(d + 64, desc_at(dl, ks), desc_at(dw, ks), idesc64, 1u);
}
// MMA issue in warp-uniform control flow.  Under a plain `threadIdx.x == 0` the compiler cannot keep the descriptors
// in uniform registers and wraps every UTCHMMA in an ELECT / R2UR / branch loop (about 13 extra instructions and two
// register-to-uniform moves per MMA on the serial issue path).  Instead: warp 0 enters (its index comes from a
// shuffle, hence provably warp-uniform), reads the operand addresses through `uni` (lane-0 broadcasts, uniform by
// construction) and one thread picked by elect.sync issues.
__device__ __forceinline__ bool issuer_warp() { return __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0) == 0; }
__device__ __forceinline__ uint32_t uni(uint32_t x) { return __shfl_sync(0xffffffffu, x, 0); }      // whole warp, inside issuer_warp()
__device__ __forceinline__ bool elect_one() {                                                      // whole warp
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ bool mma_issuer() { return issuer_warp() && elect_one(); }               // exactly one thread of warp 0
__device__ __forceinline__ void mma_commit_saddr(uint32_t bar_saddr) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_saddr) : "memory");
}
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// make the operand stores of all threads visible to the tensor core; CTA barrier
__device__ __forceinline__ void publish() {
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
}

// ---- per-CTA context --------------------------------------------------------------------------------
struct Ctx {
    uint32_t a_saddr;      // shared address of the activation operand (hi part; lo at + A_PART)
    uint32_t tmem;         // TMEM base of this CTA's 128 columns
    uint64_t* bar;         // MMA completion barrier
    uint32_t phase;
    int row, half;         // epilogue ownership of this thread
};

// CTA prologue: barrier init, TMEM allocation.  Ends with a __syncthreads().
__device__ __forceinline__ void ctx_init(Ctx& c, const Smem& s) {
    const int tid = threadIdx.x, warp = tid >> 5;
    if (tid == 0) { mbar_init(&s.bars[MMA_BAR], 1); mbar_fence_init(); }
    if (warp == 0) tmem_alloc(s.tmem_slot);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    c.a_saddr = smem_u32(s.sA);
    c.tmem = *s.tmem_slot;
    c.bar = &s.bars[MMA_BAR];
    c.phase = 0;
    c.row = (warp & 3) * 32 + (tid & 31);
    c.half = warp >> 2;
}
__device__ __forceinline__ void ctx_fini(const Ctx& c) {
    fence_before_sync();
    __syncthreads();
    if ((threadIdx.x >> 5) == 0) tmem_dealloc(c.tmem);
}

// Publish the activation operand the threads just wrote (generic proxy -> async proxy), then thread 0 issues
// the three products against the half blob at `slot` (already acquired from the ring).  Contains the CTA
// barrier that separates this stage from the previous one.
__device__ __forceinline__ void gemm_issue(const Ctx& c, const float* slot) {
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    if (issuer_warp()) {
        const uint32_t a0 = uni(c.a_saddr), b0 = uni(smem_u32(slot)), d0 = uni(c.tmem), bar = uni(smem_u32(c.bar));
        if (elect_one()) {
            // instruction descriptors: D f32, A/B f16, both K-major, M = 128; Ahi . [Whi | Wlo] as one N = 128 product (see mma_set)
            const uint32_t idesc128 = (1u << 4) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
            const uint32_t idesc64 = (1u << 4) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
            const uint32_t da = desc_lo(a0), dl = desc_lo(a0 + A_PART), dw = desc_lo(b0);
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) mma_f16(d0, desc_at(da, ks), desc_at(dw, ks), idesc128, ks > 0);
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) mma_f16(d0 + 64, desc_at(dl, ks), desc_at(dw, ks), idesc64, 1);
            mma_commit_saddr(bar);
        }
    }
}
// Wait for the products and read this thread's 32 accumulator columns: v = D1 + 2^-11 D2
__device__ __forceinline__ void gemm_result(Ctx& c, float (&v)[32]) {
    mbar_wait(c.bar, c.phase);
    c.phase ^= 1;
    fence_after_sync();
    uint32_t v1[32], v2[32];
    const uint32_t taddr = c.tmem + ((uint32_t)(c.row & ~31) << 16) + c.half * 32;
    TC_LD32(v1, taddr);
    TC_LD32(v2, taddr + 64);
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = fmaf(__uint_as_float(v2[i]), 1.0f / 2048.0f, __uint_as_float(v1[i]));
}

// ---- operand stores -----------------------------------------------------------------------------------
__device__ __forceinline__ void split2(float a, float b, __half2& hi, __half2& lo) {
    hi = __floats2half2_rn(a, b);
    const float2 f = __half22float2(hi);
    lo = __floats2half2_rn((a - f.x) * 2048.0f, (b - f.y) * 2048.0f);
}
// 32 consecutive columns [32*half, 32*half + 32) of row `row`
__device__ __forceinline__ void store_operand32(unsigned char* sA, int row, int half, const float (&v)[32]) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        __half2 h[4], l[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) split2(v[j * 8 + 2 * i], v[j * 8 + 2 * i + 1], h[i], l[i]);
        const uint32_t off = row * 128 + (((half * 4 + j) ^ (row & 7)) << 4);
        *reinterpret_cast<uint4*>(sA + off) = make_uint4(*reinterpret_cast<uint32_t*>(&h[0]), *reinterpret_cast<uint32_t*>(&h[1]),
                                                         *reinterpret_cast<uint32_t*>(&h[2]), *reinterpret_cast<uint32_t*>(&h[3]));
        *reinterpret_cast<uint4*>(sA + A_PART + off) = make_uint4(*reinterpret_cast<uint32_t*>(&l[0]), *reinterpret_cast<uint32_t*>(&l[1]),
                                                                  *reinterpret_cast<uint32_t*>(&l[2]), *reinterpret_cast<uint32_t*>(&l[3]));
    }
}
// 4 consecutive columns [4*l, 4*l + 4) of row `row` (the half-warp-per-row layout of the attention)
__device__ __forceinline__ void store_operand4(unsigned char* sA, int row, int l, const float4& v, int lo_off = A_PART) {
    __half2 h0, h1, l0, l1;
    split2(v.x, v.y, h0, l0);
    split2(v.z, v.w, h1, l1);
    const uint32_t off = row * 128 + ((((l >> 1) ^ (row & 7)) << 4) | ((l & 1) << 3));
    *reinterpret_cast<uint2*>(sA + off) = make_uint2(*reinterpret_cast<uint32_t*>(&h0), *reinterpret_cast<uint32_t*>(&h1));
    *reinterpret_cast<uint2*>(sA + lo_off + off) = make_uint2(*reinterpret_cast<uint32_t*>(&l0), *reinterpret_cast<uint32_t*>(&l1));
}

// ---- fp32 tile helpers (thread-per-row side) --------------------------------------------------------------
__device__ __forceinline__ void row_load32(const float* __restrict__ p, float (&v)[32]) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float4 x = *reinterpret_cast<const float4*>(p + j * 4);
        v[j * 4] = x.x; v[j * 4 + 1] = x.y; v[j * 4 + 2] = x.z; v[j * 4 + 3] = x.w;
    }
}
__device__ __forceinline__ void row_store32(float* __restrict__ p, const float (&v)[32]) {
#pragma unroll
    for (int j = 0; j < 8; ++j) *reinterpret_cast<float4*>(p + j * 4) = make_float4(v[j * 4], v[j * 4 + 1], v[j * 4 + 2], v[j * 4 + 3]);
}
// LayerNorm over a 64-column row held by two partner threads (32 columns each): ONE exchange of (sum, sum of squares)
// through sRed[TR][2] float2 and ONE __syncthreads(); var = E[x^2] - mean^2 (residual sums of O(1) rows: the cancellation
// costs a few ulp of the variance).  Two calls must be separated by a CTA barrier (every GEMM stage has one).
// Returns rstd; v becomes x_hat (before the affine map).
__device__ __forceinline__ float layernorm_pair(float (&v)[32], float* __restrict__ sRed, int row, int half) {
    float s = 0.f, q = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) { s += v[i]; q = fmaf(v[i], v[i], q); }
    reinterpret_cast<float2*>(sRed)[row * 2 + half] = make_float2(s, q);
    __syncthreads();
    const float4 a = *reinterpret_cast<const float4*>(sRed + row * 4);
    const float mean = (a.x + a.z) * (1.0f / 64.0f);
    const float rs = rsqrtf(fmaxf((a.y + a.w) * (1.0f / 64.0f) - mean * mean, 0.f) + 1e-5f);
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = (v[i] - mean) * rs;
    return rs;
}

}  // namespace tc
