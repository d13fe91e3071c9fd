// Tensor-core (tcgen05) attention over one env's tokens for the token-parallel passes (n_agent <= 64), sm_100a.
//
// SelfAttention of mat_src/mat/algorithms/mat/algorithm/ma_transformer.py:24-67.  An MMA tile holds 64 / LP envs, LP =
// 64, 32 or 16 token slots per env (the smallest that fits n_agent); rows of the 128-row tile are (env, head, query)
// triples, row = 2 LP e + LP h + q, keys are columns LP e + k, and a row's softmax only sees its own env's columns (the
// other envs' get zero probability, so the second product needs no masking).  For LP = 64, one env per tile:
//   S = A . K^T   A[row][0..63] = Q[q] masked to the 32 dims of head h (the other half zero), K rows = keys   (M 128, N 64, K 64)
//   P = softmax(scale S) per row (thread-per-row on the TMEM accumulator, two threads of 32 keys each per row)
//   O = P . V     P rows as the next A operand (K = keys), V read through an MN-major descriptor (no transposition);
//                 row (h, q) keeps columns 32 h .. 32 h + 31 of the 64-wide result
// Both products use the fp16 hi + scaled-lo scheme of tc.cuh (D1 = hi.hi, D2 = hi.lo + lo.hi, D = D1 + 2^-11 D2), which
// keeps the scores and outputs at fp32-class accuracy (the parity bar is 1e-4 on log-probs).  One 256-thread CTA walks
// envs; 128 TMEM columns hold S and are reused for O; 66 KB of shared memory: three CTAs per SM overlap each other's
// load -> MMA -> softmax -> MMA -> store chains.
#include <stdlib.h>

#include "policy_fwd.h"
#include "tc.cuh"

using namespace tile;

namespace {

constexpr int ATC_NT = 256;
constexpr int KV_PART = 64 * 128;            // bytes of one part (hi or lo) of a 64-row K / V operand
static_assert(KV_PART == lay::HB_LO_BYTES, "tc::mma_set expects the lo part of the B operand at HB_LO_BYTES");
constexpr size_t ATC_SMEM = 1024 + 2 * (size_t)tc::A_PART + 2 * 2 * (size_t)KV_PART + 2 * 128 * 2 * sizeof(float) + 64;

__device__ __forceinline__ uint64_t desc_mn(uint32_t saddr) {       // MN-major SWIZZLE_128B, one 64-element MN atom
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
__device__ __forceinline__ float ex2f(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

template <bool CAUSAL, int LP>
__global__ void __launch_bounds__(ATC_NT, 3)
attn_tc_kernel(const float* __restrict__ Q, const float* __restrict__ K, const float* __restrict__ V, float* __restrict__ Y,
               float* __restrict__ lse, long n_env, int L) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* base = smem_align1024(smem_raw);
    unsigned char* sA = base;                                   // [hi | lo][128 rows][128 B]: block-diagonal Q, later P
    unsigned char* sK = sA + 2 * tc::A_PART;                    // [hi | lo][64 keys][128 B]
    unsigned char* sV = sK + 2 * KV_PART;
    float* sMax = reinterpret_cast<float*>(sV + 2 * KV_PART);   // [128 rows][2 halves]
    float* sSum = sMax + 256;
    uint64_t* bar = reinterpret_cast<uint64_t*>(sSum + 256);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int tid = threadIdx.x, warp = tid >> 5;
    const int row = (warp & 3) * 32 + (tid & 31), half = warp >> 2;     // TMEM lane of this thread, key / column half
    constexpr int EPT = 64 / LP;                                        // envs per tile
    const int e_row = row / (2 * LP), h = (row / LP) & 1, q = row % LP; // this row's env within the tile, head, query
    const long n_tiles = (n_env + EPT - 1) / EPT;
    if (tid == 0) { mbar_init(bar, 1); mbar_fence_init(); }
    if (warp == 0) tc::tmem_alloc(tmem_slot);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *tmem_slot;
    const uint32_t a_addr = smem_u32(sA), k_addr = smem_u32(sK), v_addr = smem_u32(sV), bar_addr = smem_u32(bar);
    const uint32_t idesc_mn = (1u << 4) | (1u << 16) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);   // A K-major, B MN-major
    const float scale = 0.17677669529663687f * 1.4426950408889634f;     // log2(e)/sqrt(32): exp2-domain scores
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    uint32_t phase = 0;

    for (long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long env0 = tile * EPT;
        const long tok0 = (env0 + e_row) * L;                               // first token of this row's env
        const bool row_live = q < L && env0 + e_row < n_env;
        {   // L2 prefetch of the next tile's Q / K / V rows (EPT * L contiguous token rows each, two 128-byte lines per row)
            const long nt0 = (tile + gridDim.x) * EPT * (long)L, n_tok = n_env * (long)L;
            const int lines = EPT * L * 2;
            for (int i = tid; i < 3 * lines; i += ATC_NT) {
                const int mtx = i / lines, ln = i - mtx * lines;
                const long tok = nt0 + (ln >> 1);
                const float* src = mtx == 0 ? Q : (mtx == 1 ? K : V);
                if (tok < n_tok) asm volatile("prefetch.global.L2 [%0];" ::"l"(src + tok * 64 + (ln & 1) * 32));
            }
        }
        // ---- stage Q (block-diagonal over the heads), K, V as fp16 hi / lo operands; unused token slots are zero rows ------
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int idx = tid + i * ATC_NT, slot = idx >> 4, l = idx & 15;       // token slot, float4 chunk (dims 4l .. 4l+3)
            const int e = slot / LP, t = slot % LP;
            float4 qv = z4, kv = z4, vv = z4;
            if (t < L && env0 + e < n_env) {
                const long tok = (env0 + e) * L + t;
                qv = __ldg(reinterpret_cast<const float4*>(Q + tok * 64) + l);
                kv = __ldg(reinterpret_cast<const float4*>(K + tok * 64) + l);
                vv = __ldg(reinterpret_cast<const float4*>(V + tok * 64) + l);
            }
            const int hh = l >> 3;                                              // head of these dims
            tc::store_operand4(sA, e * 2 * LP + hh * LP + t, l, qv);
            tc::store_operand4(sA, e * 2 * LP + (1 - hh) * LP + t, l, z4);
            tc::store_operand4(sK, slot, l, kv, KV_PART);
            tc::store_operand4(sV, slot, l, vv, KV_PART);
        }
        tc::publish();
        // ---- S = A . K^T -----------------------------------------------------------------------------------------
        if (tc::issuer_warp()) {
            const uint32_t au = tc::uni(a_addr), ku = tc::uni(k_addr), du = tc::uni(tmem), bu = tc::uni(bar_addr);
            if (tc::elect_one()) {
                tc::mma_set(du, au, ku, false);
                tc::mma_commit_saddr(bu);
            }
        }
        mbar_wait(bar, phase);
        phase ^= 1;
        tc::fence_after_sync();
        // ---- softmax of this row over keys [32 half, 32 half + 32) with the partner thread -----------------------------
        float p[32];
        {
            const uint32_t taddr = tmem + ((uint32_t)(row & ~31) << 16) + half * 32;
            const int k0 = e_row * LP, k1 = k0 + (CAUSAL ? min(q + 1, L) : L);    // visible columns: this env's first keys
            float m = -INFINITY;
#pragma unroll
            for (int c = 0; c < 2; ++c) {                               // 16 accumulator columns at a time (register budget)
                uint32_t v1[16], v2[16];
                TC_LD16(v1, taddr + c * 16);
                TC_LD16(v2, taddr + 64 + c * 16);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const float s = fmaf(__uint_as_float(v2[i]), 1.0f / 2048.0f, __uint_as_float(v1[i])) * scale;
                    const int col = half * 32 + c * 16 + i;
                    p[c * 16 + i] = (col >= k0 && col < k1) ? s : -INFINITY;
                    m = fmaxf(m, p[c * 16 + i]);
                }
            }
            sMax[row * 2 + half] = m;
            __syncthreads();
            m = fmaxf(sMax[row * 2], sMax[row * 2 + 1]);              // the env's key 0 is always visible: finite
            float sum = 0.f;
#pragma unroll
            for (int i = 0; i < 32; ++i) { p[i] = ex2f(p[i] - m); sum += p[i]; }
            sSum[row * 2 + half] = sum;
            __syncthreads();
            sum = sSum[row * 2] + sSum[row * 2 + 1];
            const float inv = 1.0f / sum;
#pragma unroll
            for (int i = 0; i < 32; ++i) p[i] *= inv;
            if (lse && half == 0 && row_live) lse[(tok0 + q) * 2 + h] = m * 0.6931471805599453f + logf(sum);
        }
        tc::store_operand32(sA, row, half, p);                          // the S products are complete: A is free
        tc::publish();
        // ---- O = P . V (V through an MN-major descriptor: rows = keys = K index) ------------------------------------------
        if (tc::issuer_warp()) {
            const uint32_t au = tc::uni(a_addr), vu = tc::uni(v_addr), du = tc::uni(tmem), bu = tc::uni(bar_addr);
            if (tc::elect_one()) {
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {                        // 16 keys per step: A + 32 B, B + two 8-row atoms
                    const uint32_t acc = ks > 0 ? 1u : 0u;
                    const uint64_t ah = tc::make_desc(au + ks * 32), al = tc::make_desc(au + tc::A_PART + ks * 32);
                    const uint64_t bh = desc_mn(vu + ks * 2048), bl = desc_mn(vu + KV_PART + ks * 2048);
                    tc::mma_f16(du, ah, bh, idesc_mn, acc);
                    tc::mma_f16(du + 64, ah, bl, idesc_mn, acc);
                    tc::mma_f16(du + 64, al, bh, idesc_mn, 1u);
                }
                tc::mma_commit_saddr(bu);
            }
        }
        mbar_wait(bar, phase);
        phase ^= 1;
        tc::fence_after_sync();
        // ---- a row of head h keeps columns 32 h .. 32 h + 31: the thread whose column half equals its head stores them --------
        {   // the TMEM loads are warp-collective: every thread loads its half, the store is what is predicated
            const uint32_t taddr = tmem + ((uint32_t)(row & ~31) << 16) + half * 32;
            float* y = Y + (tok0 + q) * 64 + h * 32;
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                uint32_t v1[16], v2[16];
                TC_LD16(v1, taddr + c * 16);
                TC_LD16(v2, taddr + 64 + c * 16);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (row_live && half == h) {
                    float o[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) o[i] = fmaf(__uint_as_float(v2[i]), 1.0f / 2048.0f, __uint_as_float(v1[i]));
#pragma unroll
                    for (int j = 0; j < 2; ++j)                  // whole 32-byte sectors (thread-per-row: 128 bytes of one token row)
                        st_global_256(y + c * 16 + j * 8, o[j * 8], o[j * 8 + 1], o[j * 8 + 2], o[j * 8 + 3], o[j * 8 + 4], o[j * 8 + 5],
                                      o[j * 8 + 6], o[j * 8 + 7]);
                }
            }
        }
        // the next tile's staging overwrites A / K / V (their MMAs are complete) and, after its publish barrier, the accumulator
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem);
}


// ---------------------------------------------------------------------------------------------------------------------
// 64 < n_agent <= 256: one env per CTA pass, all of its K and V tiles (64 keys each) staged once; per query tile of 64
// queries (128 rows = 2 heads) the scores against every visible key tile go to their own 128 TMEM columns, the row
// maximum comes from a first pass over them, and a second pass turns one key tile at a time into probabilities (A
// operand) and accumulates O += P_kt . V_kt in the columns the first score tile occupied (it has been consumed by then).
// O is normalised in the epilogue.  512 TMEM columns and 163 KB of shared memory: one CTA per SM.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int LONG_KT = 4;                   // key / query tiles per env at most (n_agent <= 256)
constexpr size_t ATC_LONG_SMEM = 1024 + 2 * (size_t)tc::A_PART + 2 * LONG_KT * 2 * (size_t)KV_PART + 2 * 128 * 2 * sizeof(float) + 64;

template <bool CAUSAL>
__global__ void __launch_bounds__(ATC_NT, 1)
attn_tc_long_kernel(const float* __restrict__ Q, const float* __restrict__ K, const float* __restrict__ V, float* __restrict__ Y,
                    float* __restrict__ lse, long n_env, int L) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* base = smem_align1024(smem_raw);
    unsigned char* sA = base;                                       // block-diagonal Q tile, later one P tile at a time
    unsigned char* sK = sA + 2 * tc::A_PART;                        // [key tile][hi | lo][64 keys][128 B]
    unsigned char* sV = sK + LONG_KT * 2 * KV_PART;
    float* sMax = reinterpret_cast<float*>(sV + LONG_KT * 2 * KV_PART);
    float* sSum = sMax + 256;
    uint64_t* bar = reinterpret_cast<uint64_t*>(sSum + 256);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int tid = threadIdx.x, warp = tid >> 5;
    const int row = (warp & 3) * 32 + (tid & 31), half = warp >> 2;
    const int h = row >> 6, q = row & 63;
    if (tid == 0) { mbar_init(bar, 1); mbar_fence_init(); }
    if (warp == 0) tc::tmem_alloc_n<512>(tmem_slot);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *tmem_slot;
    const uint32_t a_addr = smem_u32(sA), k_addr = smem_u32(sK), v_addr = smem_u32(sV), bar_addr = smem_u32(bar);
    const uint32_t idesc_mn = (1u << 4) | (1u << 16) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const float scale = 0.17677669529663687f * 1.4426950408889634f;
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    const int n_t = (L + 63) >> 6;                                  // key tiles = query tiles
    const uint32_t lane_base = tmem + ((uint32_t)(row & ~31) << 16) + half * 32;
    uint32_t phase = 0;
    auto wait_mma = [&]() {
        mbar_wait(bar, phase);
        phase ^= 1;
        tc::fence_after_sync();
    };

    for (long env = blockIdx.x; env < n_env; env += gridDim.x) {
        const long tok0 = env * L;
        // ---- stage every K / V tile of the env (token slots >= L are zero rows) ----------------------------------------
        for (int idx = tid; idx < n_t * 64 * 16; idx += ATC_NT) {
            const int t = idx >> 4, l = idx & 15, kt = t >> 6, r = t & 63;
            float4 kv = z4, vv = z4;
            if (t < L) {
                kv = __ldg(reinterpret_cast<const float4*>(K + (tok0 + t) * 64) + l);
                vv = __ldg(reinterpret_cast<const float4*>(V + (tok0 + t) * 64) + l);
            }
            tc::store_operand4(sK + kt * 2 * KV_PART, r, l, kv, KV_PART);
            tc::store_operand4(sV + kt * 2 * KV_PART, r, l, vv, KV_PART);
        }
        for (int qt = 0; qt < n_t; ++qt) {
            const int qg = qt * 64 + q;                             // this row's query
            const bool row_live = qg < L;
            const int nkt = CAUSAL ? qt + 1 : n_t;                  // visible key tiles
            const int nk = CAUSAL ? min(qg + 1, L) : L;             // visible keys of this row
            // ---- Q tile, block-diagonal over the heads ---------------------------------------------------------
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int idx = tid + i * ATC_NT, t = idx >> 4, l = idx & 15;
                float4 qv = z4;
                if (qt * 64 + t < L) qv = __ldg(reinterpret_cast<const float4*>(Q + (tok0 + qt * 64 + t) * 64) + l);
                const int hh = l >> 3;
                tc::store_operand4(sA, hh * 64 + t, l, qv);
                tc::store_operand4(sA, (1 - hh) * 64 + t, l, z4);
            }
            tc::publish();
            if (tc::issuer_warp()) {
                const uint32_t au = tc::uni(a_addr), ku = tc::uni(k_addr), du = tc::uni(tmem), bu = tc::uni(bar_addr);
                if (tc::elect_one()) {
                    for (int kt = 0; kt < nkt; ++kt) tc::mma_set(du + kt * 128, au, ku + kt * 2 * KV_PART, false);
                    tc::mma_commit_saddr(bu);
                }
            }
            wait_mma();
            // ---- pass 1: row maximum over every visible key ----------------------------------------------------
            float m = -INFINITY;
            for (int kt = 0; kt < nkt; ++kt) {
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    uint32_t v1[16], v2[16];
                    TC_LD16(v1, lane_base + kt * 128 + c * 16);
                    TC_LD16(v2, lane_base + kt * 128 + 64 + c * 16);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    const int k0 = kt * 64 + half * 32 + c * 16;
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const float sc = fmaf(__uint_as_float(v2[i]), 1.0f / 2048.0f, __uint_as_float(v1[i])) * scale;
                        if (k0 + i < nk) m = fmaxf(m, sc);
                    }
                }
            }
            sMax[row * 2 + half] = m;
            __syncthreads();
            m = fmaxf(sMax[row * 2], sMax[row * 2 + 1]);          // key 0 is always visible: finite
            // ---- pass 2: one key tile at a time: probabilities -> A operand, O += P_kt . V_kt ---------------------------
            float sum = 0.f;
            for (int kt = 0; kt < nkt; ++kt) {
                float p[32];
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    uint32_t v1[16], v2[16];
                    TC_LD16(v1, lane_base + kt * 128 + c * 16);
                    TC_LD16(v2, lane_base + kt * 128 + 64 + c * 16);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    const int k0 = kt * 64 + half * 32 + c * 16;
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const float sc = fmaf(__uint_as_float(v2[i]), 1.0f / 2048.0f, __uint_as_float(v1[i])) * scale;
                        p[c * 16 + i] = k0 + i < nk ? ex2f(sc - m) : 0.f;
                        sum += p[c * 16 + i];
                    }
                }
                if (kt > 0) wait_mma();                            // the previous product is done with the A operand
                tc::store_operand32(sA, row, half, p);
                tc::publish();                                     // also: every thread has read score tile 0 before O overwrites it
                if (tc::issuer_warp()) {
                    const uint32_t au = tc::uni(a_addr), vu = tc::uni(v_addr + kt * 2 * KV_PART), du = tc::uni(tmem), bu = tc::uni(bar_addr);
                    if (tc::elect_one()) {
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            const uint32_t acc = (kt > 0 || ks > 0) ? 1u : 0u;
                            const uint64_t ah = tc::make_desc(au + ks * 32), al = tc::make_desc(au + tc::A_PART + ks * 32);
                            const uint64_t bh = desc_mn(vu + ks * 2048), bl = desc_mn(vu + KV_PART + ks * 2048);
                            tc::mma_f16(du, ah, bh, idesc_mn, acc);
                            tc::mma_f16(du + 64, ah, bl, idesc_mn, acc);
                            tc::mma_f16(du + 64, al, bh, idesc_mn, 1u);
                        }
                        tc::mma_commit_saddr(bu);
                    }
                }
            }
            sSum[row * 2 + half] = sum;
            wait_mma();                                            // last product
            __syncthreads();
            sum = sSum[row * 2] + sSum[row * 2 + 1];
            const float inv = 1.0f / sum;
            if (lse && half == 0 && row_live) lse[(tok0 + qg) * 2 + h] = m * 0.6931471805599453f + logf(sum);
            {
                float* y = Y + (tok0 + qg) * 64 + h * 32;
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    uint32_t v1[16], v2[16];
                    TC_LD16(v1, lane_base + c * 16);
                    TC_LD16(v2, lane_base + 64 + c * 16);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    if (row_live && half == h) {
                        float o[16];
#pragma unroll
                        for (int i = 0; i < 16; ++i) o[i] = fmaf(__uint_as_float(v2[i]), 1.0f / 2048.0f, __uint_as_float(v1[i])) * inv;
#pragma unroll
                        for (int j = 0; j < 2; ++j)
                            st_global_256(y + c * 16 + j * 8, o[j * 8], o[j * 8 + 1], o[j * 8 + 2], o[j * 8 + 3], o[j * 8 + 4], o[j * 8 + 5],
                                          o[j * 8 + 6], o[j * 8 + 7]);
                    }
                }
            }
            // the next Q tile / env overwrites A and (after its publish barrier) the score columns
        }
        __syncthreads();                                           // K / V tiles are restaged for the next env
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc_n<512>(tmem);
}

}  // namespace

int launch_attn_tc(const float* Q, const float* K, const float* V, float* Y, float* lse, long n_env, int L, int causal,
                   int n_sm, cudaStream_t st) {
    if (L > 64) {
        ARG_CHECK(L <= 64 * LONG_KT, "attn_tc: n_agent %d > %d", L, 64 * LONG_KT);
        const unsigned grid = (unsigned)(n_env < n_sm ? n_env : n_sm);
        if (causal) {
            CUDA_TRY(cudaFuncSetAttribute(attn_tc_long_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ATC_LONG_SMEM));
            attn_tc_long_kernel<true><<<grid, ATC_NT, ATC_LONG_SMEM, st>>>(Q, K, V, Y, lse, n_env, L);
        } else {
            CUDA_TRY(cudaFuncSetAttribute(attn_tc_long_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ATC_LONG_SMEM));
            attn_tc_long_kernel<false><<<grid, ATC_NT, ATC_LONG_SMEM, st>>>(Q, K, V, Y, lse, n_env, L);
        }
        CUDA_TRY(cudaGetLastError());
        return MATAS_OK;
    }
    const int lp = L <= 16 ? 16 : (L <= 32 ? 32 : 64);
    const long n_tiles = (n_env + 64 / lp - 1) / (64 / lp), resident = 3L * n_sm;
    const unsigned grid = (unsigned)(n_tiles < resident ? n_tiles : resident);
#define ATC_LAUNCH(C, P)                                                                                               \
    do {                                                                                                               \
        CUDA_TRY(cudaFuncSetAttribute(attn_tc_kernel<C, P>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ATC_SMEM)); \
        attn_tc_kernel<C, P><<<grid, ATC_NT, ATC_SMEM, st>>>(Q, K, V, Y, lse, n_env, L);                               \
    } while (0)
    if (causal) { if (lp == 16) ATC_LAUNCH(true, 16); else if (lp == 32) ATC_LAUNCH(true, 32); else ATC_LAUNCH(true, 64); }
    else { if (lp == 16) ATC_LAUNCH(false, 16); else if (lp == 32) ATC_LAUNCH(false, 32); else ATC_LAUNCH(false, 64); }
#undef ATC_LAUNCH
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}
