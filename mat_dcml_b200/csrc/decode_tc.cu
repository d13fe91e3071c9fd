// Tensor-core (tcgen05) autoregressive decode for sm_100a (get_actions / *_autoregreesive_act,
// mat_src/mat/algorithms/utils/transformer_act.py:8-100,131-173; Decoder of ma_transformer.py:157-230).
//
// One persistent 512-thread CTA per SM owns a tile of up to 128 envs and walks the decode schedule
// position by position.  Per position and decoder block:
//     x -> operand U, self cache row        FA0|FA1: q'_0, q'_1 = folded queries       (2 MMA sets, TMEM 0-255)
//     absorbed self-attention over the cached layer inputs (CUDA cores, 8 lanes per env) -> ctx_0 -> U, ctx_1 -> V
//     FB0.ctx_0 + FB1.ctx_1 (one accumulator) + x -> LN1 -> x1 -> cross cache row
//     absorbed cross-attention (queries q2' precomputed from obs_rep)                   -> ctx_0 -> U, ctx_1 -> V
//     FC0.ctx_0 + FC1.ctx_1 + obs_rep -> LN2 -> x2 -> U     M0: GELU -> h -> V     M2 + x2 -> LN3 -> x3
// then the logits head and the action pick.  All [128 x 64].[64 x 64] products are fp16 hi + scaled-lo MMAs
// with fp32 accumulators in tensor memory (tc.cuh); epilogues are thread-per-row: warp w owns rows
// 32*(w % 4) + lane and the 16 accumulator columns of quarter w / 4.
// Row traffic of the stage path: the two cache rows a position produces leave through a shared fp32 tile as whole 256-byte
// rows (tile_rows_to_cache), the block input is parked in 64 spare tensor-memory columns for the LN1 residual, obs_rep rows
// arrive by cp.async during the attention; the weight ring is fed by the CTA's last thread (thread 0 issues the MMAs).
// Block 0 is shorter than that: its self-attention attends over token embeddings (six possible rows), so it is a
// softmax over token-code counts, and with the tables of x1_table_kernel / x1_spec_table_kernel (TABK = 1 / 2) its
// x1 rows are not computed or cached per env at all: the cross-attention gathers them from an L2-resident table by
// the env's action history, and the block starts at the cross-attention.
#include <stdlib.h>

#include <type_traits>

#include "policy_fwd.h"
#include "tc.cuh"

using namespace tile;

namespace {

constexpr int NTD = 512;             // threads per CTA
constexpr int MAXR = 128;            // env rows per tile (MMA M)
constexpr int DEPTH = 4;             // weight ring slots
constexpr int TMEM_D = 512;          // columns 0-255: accumulators; 256-319: the block input row parked for the LN1 residual
constexpr int OP_BYTES = 2 * tc::A_PART;
constexpr int ROWS16 = 32;          // tiles of at most this many envs use 16 lanes per env in the attention
constexpr int PF_ROWS = 4;          // in-loop L2 prefetch distance of the cache-row loads (8 and 16 measured slower)

constexpr int STEP_BLOBS = lay::DEC_FOLD_STEP_BLOBS - 2;     // block 0 needs no query maps (see code_counts)
__constant__ int c_tc_decode[STEP_BLOBS] = {
    lay::hb(lay::fold_blob(0, lay::FB0)),
    lay::hb(lay::fold_blob(0, lay::FB1)), lay::hb(lay::fold_blob(0, lay::FC0)), lay::hb(lay::fold_blob(0, lay::FC1)),
    lay::hb(lay::dec_blob(0, lay::DM0)), lay::hb(lay::dec_blob(0, lay::DM2)),
    lay::hb(lay::fold_blob(1, lay::FA0)), lay::hb(lay::fold_blob(1, lay::FA1)), lay::hb(lay::fold_blob(1, lay::FB0)),
    lay::hb(lay::fold_blob(1, lay::FB1)), lay::hb(lay::fold_blob(1, lay::FC0)), lay::hb(lay::fold_blob(1, lay::FC1)),
    lay::hb(lay::dec_blob(1, lay::DM0)), lay::hb(lay::dec_blob(1, lay::DM2)),
    lay::hb(lay::DEC_H0)};

// the same sequence without block 0's output maps (their result comes from the table, x1_table_kernel)
__constant__ int c_tc_decode_tab[STEP_BLOBS - 2] = {
    lay::hb(lay::fold_blob(0, lay::FC0)), lay::hb(lay::fold_blob(0, lay::FC1)),
    lay::hb(lay::dec_blob(0, lay::DM0)), lay::hb(lay::dec_blob(0, lay::DM2)),
    lay::hb(lay::fold_blob(1, lay::FA0)), lay::hb(lay::fold_blob(1, lay::FA1)), lay::hb(lay::fold_blob(1, lay::FB0)),
    lay::hb(lay::fold_blob(1, lay::FB1)), lay::hb(lay::fold_blob(1, lay::FC0)), lay::hb(lay::fold_blob(1, lay::FC1)),
    lay::hb(lay::dec_blob(1, lay::DM0)), lay::hb(lay::dec_blob(1, lay::DM2)),
    lay::hb(lay::DEC_H0)};

struct DSmem {
    unsigned char* U;      // operand buffers [hi | lo][128][128 B]
    unsigned char* V;
    float* ring;           // DEPTH slots
    float* sQ0;            // [128][XS] folded self-attention queries, head 0 / head 1
    float* sQ1;
    float* sEmb;           // [6][64] token embedding table
    float* sHead;          // DEC_H3 group
    float* sRed;           // [2][128][4] LayerNorm partials
    float* sLog;           // [128][4][2] logit partials
    float* sS;             // [2 heads][6 query codes][6 key codes] block-0 self-attention scores (exp2 domain)
    int* sCode;            // [128]
    uint32_t* bits;        // [128][8] per-env action bitset
    uint64_t* bars;        // [0, DEPTH): ring, [7]: MMA completion
    uint32_t* tmem_slot;
};
constexpr size_t DSMEM_FLOATS = 384 + 160 + 1024 + 1024 + 96 + 128 + 1024;
__host__ __device__ constexpr size_t dsmem_bytes() {
    return 1024 + 2 * (size_t)OP_BYTES + (size_t)DEPTH * tc::SLOT_BYTES + 2 * (size_t)tc::TILE_BYTES + DSMEM_FLOATS * 4 + 8 * 8 + 16;
}
__device__ inline DSmem dcarve(unsigned char* raw) {
    DSmem s;
    unsigned char* p = smem_align1024(raw);
    s.U = p; p += OP_BYTES;
    s.V = p; p += OP_BYTES;
    s.ring = reinterpret_cast<float*>(p); p += (size_t)DEPTH * tc::SLOT_BYTES;
    s.sQ0 = reinterpret_cast<float*>(p); p += tc::TILE_BYTES;
    s.sQ1 = reinterpret_cast<float*>(p); p += tc::TILE_BYTES;
    float* f = reinterpret_cast<float*>(p);
    s.sEmb = f; f += 384;
    s.sHead = f; f += 160;
    s.sRed = f; f += 1024;
    s.sLog = f; f += 1024;
    s.sS = f; f += 96;
    s.sCode = reinterpret_cast<int*>(f); f += 128;
    s.bits = reinterpret_cast<uint32_t*>(f); f += 1024;
    s.bars = reinterpret_cast<uint64_t*>(f);
    s.tmem_slot = reinterpret_cast<uint32_t*>(s.bars + 8);
    return s;
}

// ---- 16-column (quarter-row) helpers ---------------------------------------------------------------
// v = D1 + 2^-11 D2 for this thread's row and quarter; accumulator set at TMEM column `cols`
__device__ __forceinline__ void result16(uint32_t tmem, int cols, int row, int q, float (&v)[16]) {
    uint32_t v1[16], v2[16];
    const uint32_t taddr = tmem + ((uint32_t)(row & ~31) << 16) + cols + q * 16;
    TC_LD16(v1, taddr);
    TC_LD16(v2, taddr + 64);
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = fmaf(__uint_as_float(v2[i]), 1.0f / 2048.0f, __uint_as_float(v1[i]));
}
// park / fetch this thread's 16 fp32 values in its own TMEM lane (columns cols + 16 q ..): a residual that survives the
// stages in between without registers, shared memory or a scattered global reload
__device__ __forceinline__ void tmem_park16(uint32_t tmem, int cols, int row, int q, const float (&v)[16]) {
    uint32_t u[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) u[i] = __float_as_uint(v[i]);
    const uint32_t taddr = tmem + ((uint32_t)(row & ~31) << 16) + cols + q * 16;
    TC_ST16(taddr, u);
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_fetch16(uint32_t tmem, int cols, int row, int q, float (&v)[16]) {
    uint32_t u[16];
    const uint32_t taddr = tmem + ((uint32_t)(row & ~31) << 16) + cols + q * 16;
    TC_LD16(u, taddr);
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(u[i]);
}
// token-major global rows (row r at src + r * row_stride floats) -> shared fp32 tile, asynchronously (no registers held):
// whole 256-byte rows per half-warp.  Complete after cp_async_wait_all() + a CTA barrier.
__device__ __forceinline__ void rows_to_tile_async(float* sT, const float* __restrict__ src, long row_stride, int n_valid) {
    for (int idx = threadIdx.x; idx < n_valid * 16; idx += NTD) {
        const int r = idx >> 4, c = idx & 15;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(sT + r * XS + c * 4)), "l"(src + (long)r * row_stride + c * 4) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void add16(float (&v)[16], const float* __restrict__ p) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float4 x = *reinterpret_cast<const float4*>(p + j * 4);
        v[j * 4] += x.x; v[j * 4 + 1] += x.y; v[j * 4 + 2] += x.z; v[j * 4 + 3] += x.w;
    }
}
__device__ __forceinline__ void load16(float (&v)[16], const float* __restrict__ p) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float4 x = *reinterpret_cast<const float4*>(p + j * 4);
        v[j * 4] = x.x; v[j * 4 + 1] = x.y; v[j * 4 + 2] = x.z; v[j * 4 + 3] = x.w;
    }
}
__device__ __forceinline__ void store16(float* __restrict__ p, const float (&v)[16]) {
#pragma unroll
    for (int j = 0; j < 4; ++j) *reinterpret_cast<float4*>(p + j * 4) = make_float4(v[j * 4], v[j * 4 + 1], v[j * 4 + 2], v[j * 4 + 3]);
}
__device__ __forceinline__ void affine16(float (&v)[16], const float* __restrict__ blob, int q) {
    const float4* w = reinterpret_cast<const float4*>(blob + lay::B_LNW + q * 16);
    const float4* b = reinterpret_cast<const float4*>(blob + lay::B_LNB + q * 16);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float4 ww = w[j], bb = b[j];
        v[j * 4] = fmaf(v[j * 4], ww.x, bb.x); v[j * 4 + 1] = fmaf(v[j * 4 + 1], ww.y, bb.y);
        v[j * 4 + 2] = fmaf(v[j * 4 + 2], ww.z, bb.z); v[j * 4 + 3] = fmaf(v[j * 4 + 3], ww.w, bb.w);
    }
}
// columns [16q, 16q + 16) of row `row` as fp16 hi / scaled-lo operand chunks
__device__ __forceinline__ void store_operand16(unsigned char* op, int row, int q, const float (&v)[16]) {
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        __half2 h[4], l[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) tc::split2(v[j * 8 + 2 * i], v[j * 8 + 2 * i + 1], h[i], l[i]);
        const uint32_t off = row * 128 + (((q * 2 + j) ^ (row & 7)) << 4);
        *reinterpret_cast<uint4*>(op + off) = make_uint4(*reinterpret_cast<uint32_t*>(&h[0]), *reinterpret_cast<uint32_t*>(&h[1]),
                                                         *reinterpret_cast<uint32_t*>(&h[2]), *reinterpret_cast<uint32_t*>(&h[3]));
        *reinterpret_cast<uint4*>(op + tc::A_PART + off) = make_uint4(*reinterpret_cast<uint32_t*>(&l[0]), *reinterpret_cast<uint32_t*>(&l[1]),
                                                                      *reinterpret_cast<uint32_t*>(&l[2]), *reinterpret_cast<uint32_t*>(&l[3]));
    }
}
// the same columns back as fp32: hi + 2^-11 lo (equal to the stored value to 2^-22 relative)
__device__ __forceinline__ void add_operand16(float (&v)[16], const unsigned char* op, int row, int q) {
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const uint32_t off = row * 128 + (((q * 2 + j) ^ (row & 7)) << 4);
        const uint4 h = *reinterpret_cast<const uint4*>(op + off);
        const uint4 l = *reinterpret_cast<const uint4*>(op + tc::A_PART + off);
        const uint32_t hw[4] = {h.x, h.y, h.z, h.w}, lw[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float2 hf = __half22float2(*reinterpret_cast<const __half2*>(&hw[i]));
            const float2 lf = __half22float2(*reinterpret_cast<const __half2*>(&lw[i]));
            v[j * 8 + 2 * i] += fmaf(lf.x, 1.0f / 2048.0f, hf.x);
            v[j * 8 + 2 * i + 1] += fmaf(lf.y, 1.0f / 2048.0f, hf.y);
        }
    }
}
// Cache rows leave through a shared fp32 tile.  In the epilogue layout a warp's store instruction would write 16 bytes
// of 32 different rows (32 half-used sectors; measured ~2,000 cycles per row store); staged, each warp instruction writes
// two whole 256-byte rows.  Row r of the tile goes to dst + r * slot_stride.
__device__ __forceinline__ void tile_rows_to_cache(const float* __restrict__ sT, float* __restrict__ dst, long slot_stride, int n_valid) {
    for (int idx = threadIdx.x; idx < n_valid * 16; idx += NTD) {
        const int r = idx >> 4, c = idx & 15;
        *reinterpret_cast<float4*>(dst + (long)r * slot_stride + c * 4) = *reinterpret_cast<const float4*>(sT + r * XS + c * 4);
    }
}

// LayerNorm of a 64-column row held by four threads (16 columns each): ONE exchange of (sum, sum of squares) through
// shared memory and ONE CTA barrier; var = E[x^2] - mean^2 (inputs are residual sums of O(1) rows: the cancellation costs
// a few ulp of the variance).  v becomes x_hat.  Two calls must be separated by a CTA barrier (every stage has one).
__device__ __forceinline__ void layernorm_quad(float (&v)[16], float* __restrict__ sRed, int row, int q) {
    float s = 0.f, d = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) { s += v[i]; d = fmaf(v[i], v[i], d); }
    reinterpret_cast<float2*>(sRed)[row * 4 + q] = make_float2(s, d);
    __syncthreads();
    const float4 a = *reinterpret_cast<const float4*>(sRed + row * 8);
    const float4 c = *reinterpret_cast<const float4*>(sRed + row * 8 + 4);
    const float mean = (a.x + a.z + c.x + c.z) * (1.0f / 64.0f);
    const float var = fmaxf((a.y + a.w + c.y + c.w) * (1.0f / 64.0f) - mean * mean, 0.f);
    const float rs = rsqrtf(var + 1e-5f);
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = (v[i] - mean) * rs;
}

// Block-0 self-attention attends over TOKEN EMBEDDINGS, which take one of six values (codes: 0 zero row,
// 1 start, 2/3 previous action 0/1, 4/5 start + own action of the Discrete variant, tile::build_emb_table), so
// its softmax over positions 0..j is a softmax over codes weighted by how many visible positions carry each
// code: no cached rows, no query GEMM.  n[c] = number of positions t <= j whose cached token is c at this
// point of the decode schedule (DecSched: `s` is the first position of the current stride block).
__device__ __forceinline__ int popc_range(const uint32_t* __restrict__ bt, int lo, int hi) {   // set bits in [lo, hi)
    int n = 0;
    if (hi <= lo) return 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
        const int b0 = w * 32;
        const int a = max(lo - b0, 0), e = min(hi - b0, 32);
        if (e > a) {
            uint32_t m = (e == 32 ? 0xffffffffu : ((1u << e) - 1u)) & ~((1u << a) - 1u);
            n += __popc(bt[w] & m);
        }
    }
    return n;
}
__device__ __forceinline__ void code_counts(const uint32_t* __restrict__ bt, int mode, int det, int kind, int j, int s, int (&n)[6]) {
#pragma unroll
    for (int c = 0; c < 6; ++c) n[c] = 0;
    if (mode == MATAS_MODE_SEMI || !det) {
        // predecessor codes (1 at t = 0, 2 + a[t-1] after); a speculative pass has zero rows behind the block start
        const int ntrue = (det && kind == 1) ? s + 1 : j + 1;
        const int ones = popc_range(bt, 0, ntrue - 1);
        n[1] = 1; n[3] = ones; n[2] = ntrue - 1 - ones; n[0] = j + 1 - ntrue;
    } else {
        // Discrete deterministic decode: committed positions carry their OWN action (4/5 at t = 0, 2/3 after)
        const int nown = kind == 2 ? j + 1 : s;
        if (nown >= 1) {
            const int ones = popc_range(bt, 1, nown);
            n[4 + (int)(bt[0] & 1u)] = 1; n[3] = ones; n[2] = nown - 1 - ones;
        }
        if (kind != 2) {                       // speculative rows of the current block: start token at t = 0, else zero rows
            if (s == 0) { n[1] = 1; n[0] = j; } else n[0] = j + 1 - s;
        }
    }
}

__device__ __forceinline__ float ex2(float x) {          // 2^x, x <= 0 here; denormal results flush to zero
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// Absorbed attention (tile::attention_absorbed) for 512 threads with tensor-core operand outputs: 8 lanes
// per env, lane g holds columns 4g..4g+3 and 32+4g..35+4g of every 64-float row; folded queries from the
// shared tiles (GQ = false) or from global rows q0 / q1 + r * q_stride (GQ = true); ctx_0 -> U, ctx_1 -> V.
// Env row r attends to nk cached rows at cache + r * slot_stride; rows r >= n_valid store zero contexts.
//
// The key loop is the kernel's hot loop (three attentions x (j + 1) keys per position), so it is written for
// instruction count: `rows` is a multiple of 4 (decode_tc_rows; a warp holds four lane groups), hence whole warps enter and every loop bound is
// warp-uniform (nk), which allows full-mask shuffles; each half of a lane group keeps the head it collects as
// "own" and the other as "oth" (lanes 0-3: own = head 0, lanes 4-7: own = head 1), so no per-key selects are
// needed; batches of two keys run unpredicated while six keys remain, and a predicated tail finishes the row.
// TAB: the keys are rows of the shared block-0 table (x1_table_kernel) instead of a per-env cache: key t of env r is
// row t (t + 1) + 2 n3 + a_{t-1}, n3 = ones among the env's actions a_0 .. a_{t-1} (its bitset at bits + 8 r); the row
// indices are produced in key order by a running popcount, and nothing is prefetched (the table lives in L2).
template <bool GQ, int PF, int TAB = 0>
__device__ __forceinline__ void attention_absorbed_op(const float* __restrict__ q0, const float* __restrict__ q1, long q_stride,
                                                      const float* cache, long slot_stride, int rows, int n_valid, int nk,
                                                      unsigned char* U, unsigned char* V, const uint32_t* bits = nullptr,
        const float* spec = nullptr, int spec_s = -1) {
    constexpr int KB = 2;
    constexpr unsigned FULL = 0xffffffffu;
    const int tid = threadIdx.x, grp = tid >> 3, g = tid & 7;
    const bool up = (g & 4) != 0;
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = grp; r < rows; r += NTD / 8) {
        const bool valid = r < n_valid;
        const int rr = valid ? r : n_valid - 1;      // rows past the batch walk a live env's rows (same trip counts) and store zeros
        float4 qa_own, qb_own, qa_oth, qb_oth;
        if (GQ) {
            const float4* po = reinterpret_cast<const float4*>((up ? q1 : q0) + (long)rr * q_stride);
            const float4* pt = reinterpret_cast<const float4*>((up ? q0 : q1) + (long)rr * q_stride);
            qa_own = __ldg(po + g); qb_own = __ldg(po + 8 + g);
            qa_oth = __ldg(pt + g); qb_oth = __ldg(pt + 8 + g);
        } else {
            const float* po = (up ? q1 : q0) + rr * XS;
            const float* pt = (up ? q0 : q1) + rr * XS;
            qa_own = *reinterpret_cast<const float4*>(po + g * 4);
            qb_own = *reinterpret_cast<const float4*>(po + 32 + g * 4);
            qa_oth = *reinterpret_cast<const float4*>(pt + g * 4);
            qb_oth = *reinterpret_cast<const float4*>(pt + 32 + g * 4);
        }
        const float4* xp = reinterpret_cast<const float4*>(TAB ? cache : cache + (long)rr * slot_stride) + g;   // row t: xp[t*16], xp[t*16 + 8]
        // L2 prefetch PF rows ahead by lanes 0 (first 128-byte line of the row) and 4 (second line); no registers held
        const float4* pfp = xp + (g ? 8 : 0) + PF * 16;
        const int pf_lim = (!TAB && PF > 0 && (g & 3) == 0) ? nk - PF : -(1 << 30);       // prefetch row t + PF while t < pf_lim
        const uint32_t* bt = TAB ? bits + rr * 8 : nullptr;
        int tab_t = 0, tab_n3 = 0;                           // TAB: next key and the ones among the actions before it
        // float4 offset (relative to the standard table) of key tab_t's row, then advance (keys come in order).  In a
        // speculative pass (spec_s >= 0) the keys behind the block start are that pass's own rows: `spec` holds them by
        // (position, ones among the actions before the block start); xp and the spec rows share one pointer base.
        const long spec_off = TAB == 2 && spec_s >= 0 ? (long)((spec - cache) >> 2) : 0;
        auto tab_row = [&]() -> long {
            if (TAB == 2 && spec_s >= 0 && tab_t > spec_s) {
                const long row = (long)(tab_t - spec_s - 1) * (spec_s + 1) + tab_n3;
                ++tab_t;
                return spec_off + row * 16;
            }
            const int ap = tab_t > 0 ? (int)((bt[(tab_t - 1) >> 5] >> ((tab_t - 1) & 31)) & 1u) : 0;
            tab_n3 += ap;
            const long row = (long)tab_t * (tab_t + 1) + 2 * tab_n3 + ap;
            ++tab_t;
            return row * 16;
        };
        float m_own = -INFINITY, m_oth = -INFINITY, l_own = 0.f, l_oth = 0.f;
        float4 ca_own = z4, cb_own = z4, ca_oth = z4, cb_oth = z4;
        auto batch = [&](const float4 (&xa)[KB], const float4 (&xb)[KB], int t0, auto masked) {
            float so[KB], st[KB];
#pragma unroll
            for (int i = 0; i < KB; ++i) {
                const float p_own = dot8_2(qa_own, xa[i], qb_own, xb[i]);       // packed FMAs: half the issue slots
                const float p_oth = dot8_2(qa_oth, xa[i], qb_oth, xb[i]);
                float v = p_own + __shfl_xor_sync(FULL, p_oth, 4);
                v += __shfl_xor_sync(FULL, v, 2);
                v += __shfl_xor_sync(FULL, v, 1);
                const float o = __shfl_xor_sync(FULL, v, 4);
                if (decltype(masked)::value) {
                    const bool ok = t0 + i < nk;
                    so[i] = ok ? v : -INFINITY;
                    st[i] = ok ? o : -INFINITY;
                } else {
                    so[i] = v; st[i] = o;
                }
            }
            const float n_own = fmaxf(m_own, fmaxf(so[0], so[1])), n_oth = fmaxf(m_oth, fmaxf(st[0], st[1]));
            const float a_own = ex2(m_own - n_own), a_oth = ex2(m_oth - n_oth);      // 2^(-inf) = 0 on the first batch
            l_own *= a_own; l_oth *= a_oth;
            scale4_2(ca_own, a_own); scale4_2(cb_own, a_own); scale4_2(ca_oth, a_oth); scale4_2(cb_oth, a_oth);
#pragma unroll
            for (int i = 0; i < KB; ++i) {
                const float p_own = ex2(so[i] - n_own), p_oth = ex2(st[i] - n_oth);
                l_own += p_own; l_oth += p_oth;
                axpy4_2(ca_own, p_own, xa[i]); axpy4_2(cb_own, p_own, xb[i]); axpy4_2(ca_oth, p_oth, xa[i]); axpy4_2(cb_oth, p_oth, xb[i]);
            }
            m_own = n_own; m_oth = n_oth;
        };
        auto fetch_masked = [&](float4 (&xa)[KB], float4 (&xb)[KB], int t0) {
#pragma unroll
            for (int i = 0; i < KB; ++i) {
                const bool ok = t0 + i < nk;
                if (TAB) {
                    if (ok) { const long o = tab_row(); xa[i] = xp[o]; xb[i] = xp[o + 8]; } else { xa[i] = z4; xb[i] = z4; }
                } else {
                    xa[i] = ok ? xp[(long)(t0 + i) * 16] : z4;
                    xb[i] = ok ? xp[(long)(t0 + i) * 16 + 8] : z4;
                    if (t0 + i < pf_lim) asm volatile("prefetch.global.L2 [%0];" ::"l"(pfp + (long)(t0 + i) * 16));
                }
            }
        };
        float4 xa[KB], xb[KB], ya[KB], yb[KB];
        fetch_masked(xa, xb, 0);
        int t0 = 0;
        {   // unpredicated part: keys t0 .. t0 + 5 exist
            const float4* xt = xp;
            const float4* pt = pfp;
            for (; t0 + 3 * KB <= nk; t0 += 2 * KB, xt += 2 * KB * 16, pt += 2 * KB * 16) {
#pragma unroll
                for (int i = 0; i < KB; ++i) {
                    if (TAB) { const long o = tab_row(); ya[i] = xp[o]; yb[i] = xp[o + 8]; }
                    else { ya[i] = xt[(KB + i) * 16]; yb[i] = xt[(KB + i) * 16 + 8]; }
                }
                if (t0 + KB < pf_lim) asm volatile("prefetch.global.L2 [%0];" ::"l"(pt + KB * 16));
                if (t0 + KB + 1 < pf_lim) asm volatile("prefetch.global.L2 [%0];" ::"l"(pt + (KB + 1) * 16));
                batch(xa, xb, t0, std::false_type{});
#pragma unroll
                for (int i = 0; i < KB; ++i) {
                    if (TAB) { const long o = tab_row(); xa[i] = xp[o]; xb[i] = xp[o + 8]; }
                    else { xa[i] = xt[(2 * KB + i) * 16]; xb[i] = xt[(2 * KB + i) * 16 + 8]; }
                }
                if (t0 + 2 * KB < pf_lim) asm volatile("prefetch.global.L2 [%0];" ::"l"(pt + 2 * KB * 16));
                if (t0 + 2 * KB + 1 < pf_lim) asm volatile("prefetch.global.L2 [%0];" ::"l"(pt + (2 * KB + 1) * 16));
                batch(ya, yb, t0 + KB, std::false_type{});
            }
        }
        for (; t0 < nk; t0 += 2 * KB) {          // predicated tail (at most 5 keys)
            fetch_masked(ya, yb, t0 + KB);
            batch(xa, xb, t0, std::true_type{});
            fetch_masked(xa, xb, t0 + 2 * KB);
            if (t0 + KB < nk) batch(ya, yb, t0 + KB, std::true_type{});
        }
        const float i_own = valid && nk > 0 ? 1.0f / l_own : 0.f, i_oth = valid && nk > 0 ? 1.0f / l_oth : 0.f;
        if (!valid) { ca_own = z4; cb_own = z4; ca_oth = z4; cb_oth = z4; }       // NaN-safe: never multiply walked garbage
        unsigned char* o_own = up ? V : U;
        unsigned char* o_oth = up ? U : V;
        tc::store_operand4(o_own, r, g, make_float4(ca_own.x * i_own, ca_own.y * i_own, ca_own.z * i_own, ca_own.w * i_own));
        tc::store_operand4(o_own, r, 8 + g, make_float4(cb_own.x * i_own, cb_own.y * i_own, cb_own.z * i_own, cb_own.w * i_own));
        tc::store_operand4(o_oth, r, g, make_float4(ca_oth.x * i_oth, ca_oth.y * i_oth, ca_oth.z * i_oth, ca_oth.w * i_oth));
        tc::store_operand4(o_oth, r, 8 + g, make_float4(cb_oth.x * i_oth, cb_oth.y * i_oth, cb_oth.z * i_oth, cb_oth.w * i_oth));
    }
}

// The same attention with SIXTEEN lanes per env (lane l holds columns 4l..4l+3 of every row) for tiles of at most
// 32 envs, where 8 lanes per env would leave half of the CTA's threads without work (small batches, large L).
// Four keys per batch, the next batch in flight in registers; written like the 8-lane version: warp-uniform trip
// counts (two lane groups per warp, `rows` even), full-mask shuffles, lanes 0-7 collect head 0 and lanes 8-15 head 1
// ("own"), unpredicated batches while twelve keys remain and a predicated tail.
template <bool GQ, int PF, int TAB = 0>
__device__ __forceinline__ void attention_absorbed_op16(const float* __restrict__ q0, const float* __restrict__ q1, long q_stride,
                                                        const float* cache, long slot_stride, int rows, int n_valid, int nk,
                                                        unsigned char* U, unsigned char* V, const uint32_t* bits = nullptr,
        const float* spec = nullptr, int spec_s = -1) {
    constexpr int KB = 4;
    constexpr unsigned FULL = 0xffffffffu;
    const int tid = threadIdx.x, grp = tid >> 4, l = tid & 15;
    const bool up = (l & 8) != 0;
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = grp; r < rows; r += NTD / 16) {
        const bool valid = r < n_valid;
        const int rr = valid ? r : n_valid - 1;      // rows past the batch walk a live env's rows (same trip counts) and store zeros
        float4 q_own, q_oth;
        if (GQ) {
            q_own = __ldg(reinterpret_cast<const float4*>((up ? q1 : q0) + (long)rr * q_stride) + l);
            q_oth = __ldg(reinterpret_cast<const float4*>((up ? q0 : q1) + (long)rr * q_stride) + l);
        } else {
            q_own = *reinterpret_cast<const float4*>((up ? q1 : q0) + rr * XS + l * 4);
            q_oth = *reinterpret_cast<const float4*>((up ? q0 : q1) + rr * XS + l * 4);
        }
        const float4* xp = reinterpret_cast<const float4*>(TAB ? cache : cache + (long)rr * slot_stride) + l;     // row t: xp[t * 16]
        const int pf_lim = (!TAB && PF > 0 && (l & 7) == 0) ? nk - 2 * PF : -(1 << 30);          // prefetch row t + 2 PF while t < pf_lim
        const uint32_t* bt = TAB ? bits + rr * 8 : nullptr;
        int tab_t = 0, tab_n3 = 0;
        const long spec_off = TAB == 2 && spec_s >= 0 ? (long)((spec - cache) >> 2) : 0;
        auto tab_row = [&]() -> long {                        // see attention_absorbed_op
            if (TAB == 2 && spec_s >= 0 && tab_t > spec_s) {
                const long row = (long)(tab_t - spec_s - 1) * (spec_s + 1) + tab_n3;
                ++tab_t;
                return spec_off + row * 16;
            }
            const int ap = tab_t > 0 ? (int)((bt[(tab_t - 1) >> 5] >> ((tab_t - 1) & 31)) & 1u) : 0;
            tab_n3 += ap;
            const long row = (long)tab_t * (tab_t + 1) + 2 * tab_n3 + ap;
            ++tab_t;
            return row * 16;
        };
        const float4* pfp = xp + 2 * PF * 16;
        float m_own = -INFINITY, m_oth = -INFINITY, l_own = 0.f, l_oth = 0.f;
        float4 c_own = z4, c_oth = z4;
        auto batch = [&](const float4 (&x)[KB], int t0, auto masked) {
            float so[KB], st[KB];
#pragma unroll
            for (int i = 0; i < KB; ++i) {
                const float p_own = q_own.x * x[i].x + q_own.y * x[i].y + q_own.z * x[i].z + q_own.w * x[i].w;
                const float p_oth = q_oth.x * x[i].x + q_oth.y * x[i].y + q_oth.z * x[i].z + q_oth.w * x[i].w;
                float v = p_own + __shfl_xor_sync(FULL, p_oth, 8);
                v += __shfl_xor_sync(FULL, v, 4);
                v += __shfl_xor_sync(FULL, v, 2);
                v += __shfl_xor_sync(FULL, v, 1);
                const float o = __shfl_xor_sync(FULL, v, 8);
                if (decltype(masked)::value) {
                    const bool ok = t0 + i < nk;
                    so[i] = ok ? v : -INFINITY;
                    st[i] = ok ? o : -INFINITY;
                } else {
                    so[i] = v; st[i] = o;
                }
            }
            float n_own = m_own, n_oth = m_oth;
#pragma unroll
            for (int i = 0; i < KB; ++i) { n_own = fmaxf(n_own, so[i]); n_oth = fmaxf(n_oth, st[i]); }
            const float a_own = ex2(m_own - n_own), a_oth = ex2(m_oth - n_oth);
            l_own *= a_own; l_oth *= a_oth;
            scale4_2(c_own, a_own); scale4_2(c_oth, a_oth);
#pragma unroll
            for (int i = 0; i < KB; ++i) {
                const float p_own = ex2(so[i] - n_own), p_oth = ex2(st[i] - n_oth);
                l_own += p_own; l_oth += p_oth;
                axpy4_2(c_own, p_own, x[i]); axpy4_2(c_oth, p_oth, x[i]);
            }
            m_own = n_own; m_oth = n_oth;
        };
        auto fetch_masked = [&](float4 (&x)[KB], int t0) {
#pragma unroll
            for (int i = 0; i < KB; ++i) {
                if (TAB) {
                    if (t0 + i < nk) x[i] = xp[tab_row()]; else x[i] = z4;
                } else {
                    x[i] = t0 + i < nk ? xp[(long)(t0 + i) * 16] : z4;
                    if (t0 + i < pf_lim) asm volatile("prefetch.global.L2 [%0];" ::"l"(pfp + (long)(t0 + i) * 16));
                }
            }
        };
        float4 x[KB], y[KB];
        fetch_masked(x, 0);
        int t0 = 0;
        {   // unpredicated part: keys t0 .. t0 + 11 exist
            const float4* xt = xp;
            const float4* pt = pfp;
            for (; t0 + 3 * KB <= nk; t0 += 2 * KB, xt += 2 * KB * 16, pt += 2 * KB * 16) {
#pragma unroll
                for (int i = 0; i < KB; ++i) {
                    if (TAB) y[i] = xp[tab_row()]; else y[i] = xt[(KB + i) * 16];
                    if (t0 + KB + i < pf_lim) asm volatile("prefetch.global.L2 [%0];" ::"l"(pt + (KB + i) * 16));
                }
                batch(x, t0, std::false_type{});
#pragma unroll
                for (int i = 0; i < KB; ++i) {
                    if (TAB) x[i] = xp[tab_row()]; else x[i] = xt[(2 * KB + i) * 16];
                    if (t0 + 2 * KB + i < pf_lim) asm volatile("prefetch.global.L2 [%0];" ::"l"(pt + (2 * KB + i) * 16));
                }
                batch(y, t0 + KB, std::false_type{});
            }
        }
        for (; t0 < nk; t0 += 2 * KB) {          // predicated tail (at most 11 keys)
            fetch_masked(y, t0 + KB);
            batch(x, t0, std::true_type{});
            fetch_masked(x, t0 + 2 * KB);
            if (t0 + KB < nk) batch(y, t0 + KB, std::true_type{});
        }
        const float i_own = valid && nk > 0 ? 1.0f / l_own : 0.f, i_oth = valid && nk > 0 ? 1.0f / l_oth : 0.f;
        if (!valid) { c_own = z4; c_oth = z4; }       // NaN-safe: never multiply walked garbage
        tc::store_operand4(up ? V : U, r, l, make_float4(c_own.x * i_own, c_own.y * i_own, c_own.z * i_own, c_own.w * i_own));
        tc::store_operand4(up ? U : V, r, l, make_float4(c_oth.x * i_oth, c_oth.y * i_oth, c_oth.z * i_oth, c_oth.w * i_oth));
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Block-0 cross-attention inputs as a TABLE (stochastic decode).  The row that block 0 caches for position t,
// x1_t = LN1(e_c + out-map(softmax over token codes)), depends on the env only through the code of its own token (start
// at t = 0, else the previous action a_{t-1}) and on how many of the earlier actions were 1 (code_counts): at most
// 2 (t + 1) distinct rows per position, the same for every env.  They are computed once per weight update into
// x1tab[t (t + 1) + 2 n3 + a_prev][64] (n3 = ones among a_0 .. a_{t-1}), 0.7 MB at n_agent 51, so that the decode neither
// writes nor re-reads a per-env cache for this attention (a third of its HBM traffic), skips block 0's output-map stage
// and its token-code softmax, and gathers the keys from an L2-resident table by prefix popcounts of the action bitset.
// One CTA per position; fp32 CUDA-core arithmetic (the decode's own fp16 hi/lo stage agrees to ~1e-6).
// ---------------------------------------------------------------------------------------------------------------------
struct X1Smem {
    float sEmb[6 * 64];
    float sQ[2 * 6 * 64];        // folded block-0 self-attention queries per (head, code)
    float sS[72];                // scores [head][query code][key code], exp2 domain
    float sP[2 * 6 * 64];        // P_h[c][n] = sum_k e_c[k] FB_h.Wt[k][n]
};
__device__ __forceinline__ void x1_prologue(X1Smem& m, const float* __restrict__ w) {      // 128 threads
    const int tid = threadIdx.x;
    build_emb_table(m.sEmb, w);
    __syncthreads();
    for (int idx = tid; idx < 2 * 6 * 64; idx += 128) {
        const int h = idx / 384, c = (idx / 64) % 6, n = idx & 63;
        const float* WA = w + lay::fold_blob(0, h ? lay::FA1 : lay::FA0);
        const float* WB = w + lay::fold_blob(0, h ? lay::FB1 : lay::FB0);
        float qa = WA[lay::B_BIAS + n], pb = 0.f;
        for (int k = 0; k < 64; ++k) { qa = fmaf(m.sEmb[c * 64 + k], WA[k * 64 + n], qa); pb = fmaf(m.sEmb[c * 64 + k], WB[k * 64 + n], pb); }
        m.sQ[idx] = qa; m.sP[idx] = pb;
    }
    __syncthreads();
    if (tid < 72) {
        const int h = tid / 36, cq = (tid / 6) % 6, c = tid % 6;
        float acc = 0.f;
        for (int n = 0; n < 64; ++n) acc = fmaf(m.sQ[(h * 6 + cq) * 64 + n], m.sEmb[c * 64 + n], acc);
        m.sS[tid] = acc;
    }
    __syncthreads();
}
// one warp: x1 row for own token code cq and visible-code counts cnt -> dst[64] (two columns per lane)
__device__ __forceinline__ void x1_row(const X1Smem& m, const float* __restrict__ w, int cq, const int (&cnt)[6], float* __restrict__ dst) {
    const float* W1 = w + lay::fold_blob(0, lay::FB1);                 // bias, LN weight / bias of the stage
    const int lane = threadIdx.x & 31;
    float y0 = W1[lay::B_BIAS + lane] + m.sEmb[cq * 64 + lane], y1 = W1[lay::B_BIAS + 32 + lane] + m.sEmb[cq * 64 + 32 + lane];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const float* S = m.sS + (h * 6 + cq) * 6;
        float mx = -INFINITY, wg[6], z = 0.f;
#pragma unroll
        for (int c = 0; c < 6; ++c) if (cnt[c] > 0) mx = fmaxf(mx, S[c]);
#pragma unroll
        for (int c = 0; c < 6; ++c) { wg[c] = cnt[c] > 0 ? (float)cnt[c] * exp2f(S[c] - mx) : 0.f; z += wg[c]; }
        const float iz = 1.0f / z;
#pragma unroll
        for (int c = 0; c < 6; ++c) {
            y0 = fmaf(wg[c] * iz, m.sP[(h * 6 + c) * 64 + lane], y0);
            y1 = fmaf(wg[c] * iz, m.sP[(h * 6 + c) * 64 + 32 + lane], y1);
        }
    }
    const float mean = warp_sum(y0 + y1) * (1.0f / 64.0f);
    y0 -= mean; y1 -= mean;
    const float rs = rsqrtf(warp_sum(y0 * y0 + y1 * y1) * (1.0f / 64.0f) + 1e-5f);
    dst[lane] = y0 * rs * W1[lay::B_LNW + lane] + W1[lay::B_LNB + lane];
    dst[32 + lane] = y1 * rs * W1[lay::B_LNW + 32 + lane] + W1[lay::B_LNB + 32 + lane];
}

__global__ void __launch_bounds__(128)
x1_table_kernel(const float* __restrict__ w, float* __restrict__ tab, int L) {
    __shared__ X1Smem m;
    const int t = blockIdx.x;
    x1_prologue(m, w);
    for (int e = threadIdx.x >> 5; e < 2 * (t + 1); e += 4) {          // one warp per table row
        const int n3 = e >> 1, ap = e & 1;
        if (t == 0 ? e != 0 : (ap > n3)) continue;                     // unreachable states (a_prev = 1 needs n3 >= 1)
        const int cnt[6] = {0, 1, t - n3, n3, 0, 0};
        x1_row(m, w, t == 0 ? 1 : 2 + ap, cnt, tab + ((long)t * (t + 1) + e) * 64);
    }
}

// Stride blocks [s, e) of the deterministic Semi_Discrete schedule (DecSched) and, per block, the offset of its
// speculative rows in the second table: a speculative pass writes rows t = s+1 .. e-1 with a zero-row token and the
// counts (t - s zero rows, start, s - n3 zeros, n3 ones), n3 = ones among a_0 .. a_{s-1}: (e - s - 1)(s + 1) rows
// [t - s - 1][n3] per block.
struct SpecBlocks {
    int L, nd, stride, s, e, go;
    long base;
    __host__ __device__ void init(int L_, int stride_) { L = L_; nd = L_ - 1; stride = stride_ < 1 ? 1 : stride_; s = 0; e = 1; go = 1; base = 0; }
    __host__ __device__ bool valid() const { return e <= L && go; }
    __host__ __device__ long rows() const { return (long)(e - s - 1) * (s + 1); }
    __host__ __device__ void next() {                               // mirrors DecSched::next (Semi_Discrete)
        base += rows();
        if (e == L) { go = 0; return; }
        if (e < nd) { s = e; e = min(e + stride, nd); }
        else { s = e; e = min(e + 1, L); }
    }
};

__global__ void __launch_bounds__(128)
x1_spec_table_kernel(const float* __restrict__ w, float* __restrict__ spec, int L, int stride) {
    __shared__ X1Smem m;
    const int t = blockIdx.x;
    SpecBlocks sb;
    sb.init(L, stride);
    while (sb.valid() && !(t >= sb.s && t < sb.e)) sb.next();          // the block that holds position t
    if (!sb.valid() || t == sb.s) return;                              // block starts use their true predecessor
    x1_prologue(m, w);
    const int s = sb.s;
    for (int n3 = threadIdx.x >> 5; n3 <= s; n3 += 4) {
        const int cnt[6] = {t - s, 1, s - n3, n3, 0, 0};
        x1_row(m, w, 0, cnt, spec + (sb.base + (long)(t - s - 1) * (s + 1) + n3) * 64);
    }
}

// TABK: 0 = every attention cached per env; 1 = block-0 cross-attention from the table (stochastic decode);
// 2 = the same plus the speculative rows of the deterministic stride blocks
template <int PFD, int TABK>
__global__ void __launch_bounds__(NTD, 1)
decode_tc_kernel(PolicyBufs b, DecodeArgs a, const float* __restrict__ x1tab, const float* __restrict__ x1spec) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    DSmem s = dcarve(smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5;
    const int row = (warp & 3) * 32 + (tid & 31), q = warp >> 2;
    const int L = a.L, rows = a.rows;
    constexpr int rows16 = ROWS16;
    const long n_tiles = (a.B + rows - 1) / rows;
    long my_tiles = 0;
    for (long t = blockIdx.x; t < n_tiles; t += gridDim.x) ++my_tiles;
    constexpr bool use_tab = TABK != 0;          // block 0's cross-attention keys come from the shared table(s)
    const int n_seq = use_tab ? STEP_BLOBS - 2 : STEP_BLOBS;
    Ring ring;
    // the weight ring is fed by the CTA's last thread: thread 0 issues the MMAs of a stage, and with the bulk copies on it
    // as well every stage waited for its ~700 cycles of mbarrier / copy issue
    ring.init(s.ring, s.bars, b.w, use_tab ? c_tc_decode_tab : c_tc_decode, n_seq, (int)(my_tiles * a.n_steps * n_seq), DEPTH,
              lay::BLOB_BYTES, tc::SLOT_F, NTD - 1);
    uint64_t* mma_bar = &s.bars[tc::MMA_BAR];
    if (tid == 0) { mbar_init(mma_bar, 1); mbar_fence_init(); }
    if (warp == 0) tc::tmem_alloc_n<TMEM_D>(s.tmem_slot);
    build_emb_table(s.sEmb, b.w);
    for (int i = tid; i < lay::DEC_H3_SIZE; i += NTD) s.sHead[i] = b.w[lay::DEC_H3 + i];
    for (int i = tid; i < 2 * OP_BYTES / 16; i += NTD) reinterpret_cast<uint4*>(s.U)[i] = make_uint4(0u, 0u, 0u, 0u);   // U and V
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    ring.prefetch();
    // block-0 self-attention score table: S_h[cq][c] = q'_h(e_cq) . e_c with the folded query map FA_h (fp32 blob)
    for (int idx = tid; idx < 2 * 6 * 64; idx += NTD) {
        const int h = idx / 384, cq = (idx / 64) % 6, n = idx & 63;
        const float* W = b.w + lay::fold_blob(0, h ? lay::FA1 : lay::FA0);      // Wt[k][n], bias at B_BIAS
        float acc = W[lay::B_BIAS + n];
        for (int k = 0; k < 64; ++k) acc = fmaf(s.sEmb[cq * 64 + k], W[k * 64 + n], acc);
        s.sQ0[idx] = acc;
    }
    __syncthreads();
    if (tid < 72) {
        const int h = tid / 36, cq = (tid / 6) % 6, c = tid % 6;
        float acc = 0.f;
        for (int n = 0; n < 64; ++n) acc = fmaf(s.sQ0[(h * 6 + cq) * 64 + n], s.sEmb[c * 64 + n], acc);
        s.sS[tid] = acc;
    }
    __syncthreads();
    const uint32_t tmem = *s.tmem_slot;
    const uint32_t uaddr = smem_u32(s.U), vaddr = smem_u32(s.V);
    uint32_t phase = 0;
    auto mma_wait = [&]() {
        mbar_wait(mma_bar, phase);
        phase ^= 1;
        tc::fence_after_sync();
    };

    const long slot_stride = 4L * L * 64;      // [blk 2][self | cross][L][64] cached attention inputs per env slot
    float* kv0 = b.kv + (long)blockIdx.x * rows * slot_stride;
    const int nd = a.mode == MATAS_MODE_SEMI ? L - 1 : L;
    // L2 prefetch of one token-major row per live env (two lines each) a phase before its first use: the folded
    // cross-attention queries and the obs_rep rows are otherwise first touched at the head of a dependency chain
    auto prefetch_env_rows = [&](const float* row0, int n_live) {
        if (tid < 2 * n_live) asm volatile("prefetch.global.L2 [%0];" ::"l"(row0 + (long)(tid >> 1) * L * 64 + (tid & 1) * 32));
    };
    const float sd1 = 0.5f / (1.0f + expf(-s.sHead[lay::DEC_H3_LOGSTD + 1]));
    float* sQ0row = s.sQ0 + row * XS + q * 16;
    float* sQ1row = s.sQ1 + row * XS + q * 16;

    for (long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long env0 = tile * rows;
        const int n_valid = (int)min((long)rows, a.B - env0);
        const bool live = row < n_valid;           // this thread's row maps to an env
        for (int i = tid; i < MAXR * 8; i += NTD) s.bits[i] = 0u;
        DecSched sched;
        sched.init(L, a.mode, a.deterministic, a.stride);
        DecStep st;
        SpecBlocks sb;                             // table mode, deterministic decode: the current stride block and its rows
        sb.init(L, a.stride);
        __syncthreads();
        while (sched.next(st)) {
            const int j = st.j;
            // availability of (env = tid, agent j) for the pick at the end of the position: loaded a position's worth of
            // stages ahead of its use
            float2 av = make_float2(1.f, 1.f);
            if (a.ava && tid < n_valid) av = __ldg(reinterpret_cast<const float2*>(a.ava + ((env0 + tid) * L + j) * 2));
            if (TABK == 2) while (sb.s != sched.s) sb.next();
            // ---- input token of this position (with the table, block 0 never looks at it) -----------------------
            float v[16];
            if (!use_tab) {
                if (tid < MAXR) {
                    int code;
                    const uint32_t* bt = s.bits + tid * 8;
                    if (st.kind == 1) code = j == 0 ? 1 : 0;
                    else if (st.kind == 0) code = j == 0 ? 1 : 2 + (int)((bt[(j - 1) >> 5] >> ((j - 1) & 31)) & 1u);
                    else code = (j == 0 ? 4 : 2) + (int)((bt[j >> 5] >> (j & 31)) & 1u);
                    s.sCode[tid] = code;
                }
                __syncthreads();
                load16(v, s.sEmb + s.sCode[row] * 64 + q * 16);
            }
            for (int blk = 0; blk < 2; ++blk) {
                float* cx1 = kv0 + (blk * 2 + 0) * (long)L * 64;     // self-attention inputs (block input x)
                float* cx2 = kv0 + (blk * 2 + 1) * (long)L * 64;     // cross-attention inputs (x after LN1)
                if (blk == 0 && use_tab) {
                    // ---- block 0 from the table: no token-code softmax, no output-map stage, no cache row; the
                    //      cross-attention gathers its keys x1_t from x1tab by the env's action history ------------------
                    // keys behind the block start in a speculative pass (kind 1) are that pass's rows (second table)
                    const int spec_s = (TABK == 2 && st.kind == 1) ? sched.s : -1;
                    const float* spec = x1spec + sb.base * 64;
                    rows_to_tile_async(s.sQ1, b.rep + (env0 * L + j) * 64, (long)L * 64, n_valid);       // LN2's residual rows
                    if (rows <= rows16)
                        attention_absorbed_op16<true, PFD, TABK>(b.Q2p[0][0] + (env0 * L + j) * 64, b.Q2p[0][1] + (env0 * L + j) * 64, (long)L * 64,
                                                                 x1tab, 0, rows, n_valid, j + 1, s.U, s.V, s.bits, spec, spec_s);
                    else
                        attention_absorbed_op<true, PFD, TABK>(b.Q2p[0][0] + (env0 * L + j) * 64, b.Q2p[0][1] + (env0 * L + j) * 64, (long)L * 64,
                                                               x1tab, 0, rows, n_valid, j + 1, s.U, s.V, s.bits, spec, spec_s);
                } else {
                if (blk == 0) {
                    // ---- block 0: softmax over token codes (see code_counts) -> ctx_0 -> U, ctx_1 -> V ------------
                    const int cq = s.sCode[row];
                    int cnt[6];
                    code_counts(s.bits + row * 8, a.mode, a.deterministic, st.kind, j, sched.s, cnt);
                    float wgt[2][6];
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const float* S = s.sS + (h * 6 + cq) * 6;
                        float m = -INFINITY;
#pragma unroll
                        for (int c = 0; c < 6; ++c) if (cnt[c] > 0) m = fmaxf(m, S[c]);
                        float z = 0.f;
#pragma unroll
                        for (int c = 0; c < 6; ++c) { wgt[h][c] = cnt[c] > 0 ? (float)cnt[c] * exp2f(S[c] - m) : 0.f; z += wgt[h][c]; }
                        const float iz = 1.0f / z;
#pragma unroll
                        for (int c = 0; c < 6; ++c) wgt[h][c] *= iz;
                    }
                    float c0[16], c1[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) { c0[i] = 0.f; c1[i] = 0.f; }
#pragma unroll
                    for (int c = 0; c < 6; ++c) {
                        float e[16];
                        load16(e, s.sEmb + c * 64 + q * 16);
#pragma unroll
                        for (int i = 0; i < 16; ++i) { c0[i] = fmaf(wgt[0][c], e[i], c0[i]); c1[i] = fmaf(wgt[1][c], e[i], c1[i]); }
                    }
                    store_operand16(s.U, row, q, c0);
                    store_operand16(s.V, row, q, c1);
                } else {
                // ---- folded self-attention queries ---------------------------------------------------
                store_operand16(s.U, row, q, v);
                store16(sQ1row, v);                               // the block's input row, on its way to the self cache
                tmem_park16(tmem, 256, row, q, v);                // and kept for the LN1 residual
                {
                    const float* w0 = ring.peek(0);
                    const float* w1 = ring.peek(1);
                    tc::publish();
                    if (tc::mma_issuer()) {
                        tc::mma_set(tmem, uaddr, smem_u32(w0), false);
                        tc::mma_set(tmem + 128, uaddr, smem_u32(w1), false);
                        tc::mma_commit(mma_bar);
                    }
                    ring.prefetch();
                    tile_rows_to_cache(s.sQ1, cx1 + (long)j * 64, slot_stride, n_valid);     // under the MMAs
                    __syncthreads();                              // the query tiles are rewritten below
                    mma_wait();
                    result16(tmem, 0, row, q, v);
                    add16(v, w0 + lay::B_BIAS + q * 16);
                    store16(sQ0row, v);
                    result16(tmem, 128, row, q, v);
                    add16(v, w1 + lay::B_BIAS + q * 16);
                    store16(sQ1row, v);
                    ring.release(); ring.release();
                }
                tc::fence_before_sync();
                __syncthreads();
                prefetch_env_rows(b.Q2p[1][0] + (env0 * L + j) * 64, n_valid);
                prefetch_env_rows(b.Q2p[1][1] + (env0 * L + j) * 64, n_valid);
                if (rows <= rows16) attention_absorbed_op16<false, PFD>(s.sQ0, s.sQ1, 0, cx1, slot_stride, rows, n_valid, j + 1, s.U, s.V);
                else attention_absorbed_op<false, PFD>(s.sQ0, s.sQ1, 0, cx1, slot_stride, rows, n_valid, j + 1, s.U, s.V);
                }
                // ---- self-attention output maps + x -> LN1 -> cross cache row -------------------------
                {
                    const float* w0 = ring.peek(0);
                    const float* w1 = ring.peek(1);
                    tc::publish();
                    if (tc::mma_issuer()) {
                        tc::mma_set(tmem, uaddr, smem_u32(w0), false);
                        tc::mma_set(tmem, vaddr, smem_u32(w1), true);
                        tc::mma_commit(mma_bar);
                    }
                    ring.prefetch();
                    mma_wait();
                    result16(tmem, 0, row, q, v);
                    add16(v, w1 + lay::B_BIAS + q * 16);
                    if (blk == 0) add16(v, s.sEmb + s.sCode[row] * 64 + q * 16);
                    else {
                        float res[16];                             // the block input, parked in tensor memory two stages ago
                        tmem_fetch16(tmem, 256, row, q, res);
#pragma unroll
                        for (int i = 0; i < 16; ++i) v[i] += res[i];
                    }
                    layernorm_quad(v, s.sRed, row, q);
                    affine16(v, w1, q);
                    ring.release(); ring.release();
                    store16(sQ0row, v);                            // x1 row, on its way to the cross cache (the query tiles are free)
                }
                tc::fence_before_sync();
                __syncthreads();
                tile_rows_to_cache(s.sQ0, cx2 + (long)j * 64, slot_stride, n_valid);
                rows_to_tile_async(s.sQ1, b.rep + (env0 * L + j) * 64, (long)L * 64, n_valid);           // LN2's residual rows
                __syncthreads();
                // ---- cross-attention: folded queries q2' precomputed from obs_rep ------------------------
                if (rows <= rows16)
                    attention_absorbed_op16<true, PFD>(b.Q2p[blk][0] + (env0 * L + j) * 64, b.Q2p[blk][1] + (env0 * L + j) * 64, (long)L * 64,
                                                       cx2, slot_stride, rows, n_valid, j + 1, s.U, s.V);
                else
                    attention_absorbed_op<true, PFD>(b.Q2p[blk][0] + (env0 * L + j) * 64, b.Q2p[blk][1] + (env0 * L + j) * 64, (long)L * 64,
                                                     cx2, slot_stride, rows, n_valid, j + 1, s.U, s.V);
                }
                {
                    const float* w0 = ring.peek(0);
                    const float* w1 = ring.peek(1);
                    cp_async_wait_all();                           // this thread's share of the obs_rep rows has landed in sQ1
                    tc::publish();
                    if (tc::mma_issuer()) {
                        tc::mma_set(tmem, uaddr, smem_u32(w0), false);
                        tc::mma_set(tmem, vaddr, smem_u32(w1), true);
                        tc::mma_commit(mma_bar);
                    }
                    ring.prefetch();
                    mma_wait();
                    result16(tmem, 0, row, q, v);
                    add16(v, w1 + lay::B_BIAS + q * 16);
                    if (live) add16(v, sQ1row);                    // + obs_rep (rows past the batch hold stale data)
                    layernorm_quad(v, s.sRed, row, q);
                    affine16(v, w1, q);
                    ring.release(); ring.release();
                    store_operand16(s.U, row, q, v);               // x2
                }
                // ---- MLP ---------------------------------------------------------------------------------
                {
                    const float* w0 = ring.peek(0);
                    tc::publish();
                    if (tc::mma_issuer()) { tc::mma_set(tmem, uaddr, smem_u32(w0), false); tc::mma_commit(mma_bar); }
                    ring.prefetch();
                    mma_wait();
                    result16(tmem, 0, row, q, v);
                    add16(v, w0 + lay::B_BIAS + q * 16);
                    ring.release();
#pragma unroll
                    for (int i = 0; i < 16; i += 2) gelu_erf_x2(v[i], v[i + 1]);
                    store_operand16(s.V, row, q, v);
                }
                {
                    const float* w0 = ring.peek(0);
                    tc::publish();
                    if (tc::mma_issuer()) { tc::mma_set(tmem, vaddr, smem_u32(w0), false); tc::mma_commit(mma_bar); }
                    ring.prefetch();
                    mma_wait();
                    result16(tmem, 0, row, q, v);
                    add16(v, w0 + lay::B_BIAS + q * 16);
                    add_operand16(v, s.U, row, q);                 // + x2
                    layernorm_quad(v, s.sRed, row, q);
                    affine16(v, w0, q);
                    ring.release();
                }
            }
            if (j + 1 < L) {
                prefetch_env_rows(b.Q2p[0][0] + (env0 * L + j + 1) * 64, n_valid);
                prefetch_env_rows(b.Q2p[0][1] + (env0 * L + j + 1) * 64, n_valid);
                prefetch_env_rows(b.rep + (env0 * L + j + 1) * 64, n_valid);
            }
            // ---- logits head + action pick -------------------------------------------------------------
            if (st.emit) {
                store_operand16(s.U, row, q, v);
                const float* w0 = ring.peek(0);
                tc::publish();
                if (tc::mma_issuer()) { tc::mma_set(tmem, uaddr, smem_u32(w0), false); tc::mma_commit(mma_bar); }
                ring.prefetch();
                mma_wait();
                result16(tmem, 0, row, q, v);
                add16(v, w0 + lay::B_BIAS + q * 16);
#pragma unroll
                for (int i = 0; i < 16; i += 2) gelu_erf_x2(v[i], v[i + 1]);
                layernorm_quad(v, s.sRed, row, q);
                affine16(v, w0, q);
                ring.release();
                float d0 = 0.f, d1 = 0.f;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    d0 = fmaf(v[i], s.sHead[q * 16 + i], d0);
                    d1 = fmaf(v[i], s.sHead[64 + q * 16 + i], d1);
                }
                s.sLog[(row * 4 + q) * 2] = d0;
                s.sLog[(row * 4 + q) * 2 + 1] = d1;
                tc::fence_before_sync();
                __syncthreads();
                if (tid < n_valid) {
                    const long env = env0 + tid;
                    const long g = env * L + j;
                    const float* lp = s.sLog + tid * 8;
                    const float l0 = lp[0] + lp[2] + lp[4] + lp[6] + s.sHead[lay::DEC_H3_B];
                    const float l1 = lp[1] + lp[3] + lp[5] + lp[7] + s.sHead[lay::DEC_H3_B + 1];
                    if (j < nd) {
                        float z0, z1;
                        masked_logz(l0, l1, &av.x, z0, z1);
                        int act;
                        if (a.deterministic) {
                            act = z1 > z0;
                        } else {
                            const float zm = fmaxf(z0, z1);
                            const float e0 = expf(z0 - zm), e1 = expf(z1 - zm);
                            const float p0 = e0 / (e0 + e1);
                            const float u = a.u ? a.u[g]
                                                : u01f(philox4x32((uint32_t)(a.env_offset + env), a.step, rng_tag(S_POLICY, j), 0, a.seed).x);
                            act = u >= p0;
                        }
                        a.actions[g] = (float)act;
                        a.logp[g] = act ? z1 : z0;
                        if (act) s.bits[tid * 8 + (j >> 5)] |= 1u << (j & 31);
                        else s.bits[tid * 8 + (j >> 5)] &= ~(1u << (j & 31));
                    } else {
                        float z = 0.f;
                        if (!a.deterministic) {
                            if (a.z) z = a.z[env * 2 + 1];
                            else {
                                const uint4 x = philox4x32((uint32_t)(a.env_offset + env), a.step, rng_tag(S_POLICY, j), 0, a.seed);
                                const float r1 = sqrtf(-2.0f * logf(((float)(x.z >> 8) + 1.0f) * (1.0f / 16777216.0f)));
                                z = r1 * cospif(2.0f * ((float)(x.w >> 8) * (1.0f / 16777216.0f)));
                            }
                        }
                        const float act = l1 + sd1 * z;
                        const float d = act - l1;
                        a.actions[g] = act;
                        a.logp[g] = -(d * d) / (2.0f * sd1 * sd1) - logf(sd1) - HALF_LOG_2PI;
                    }
                }
            } else {
                ring.peek(0);
                ring.release();
            }
            __syncthreads();
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc_n<TMEM_D>(tmem);
}

}  // namespace

long decode_tc_spec_rows(int L, int stride) {
    SpecBlocks sb;
    sb.init(L, stride);
    while (sb.valid()) sb.next();
    return sb.base;
}

int launch_x1_spec_table(const float* packed, float* x1spec, int L, int stride, cudaStream_t st) {
    x1_spec_table_kernel<<<L, 128, 0, st>>>(packed, x1spec, L, stride);
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}

int launch_x1_table(const float* packed, float* x1tab, int L, cudaStream_t st) {
    x1_table_kernel<<<L, 128, 0, st>>>(packed, x1tab, L);
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}

// rows per tile: spread the envs over the SMs (one CTA each), a multiple of 4 (whole warps of four 8-lane groups), at most
// 128.  4096 envs on 148 SMs: 28 rows on 147 CTAs (a multiple of 8 would put 32 rows on 128 CTAs and leave 20 SMs idle)
int decode_tc_rows(long B, int n_sm) {
    long r = (B + n_sm - 1) / n_sm;
    r = (r + 3) / 4 * 4;
    if (r < 4) r = 4;
    if (r > MAXR) r = MAXR;
    return (int)r;
}

int decode_tc_geometry(long B, int n_sm, int* grid, long* slots, int* rows_out) {
    const int rows = decode_tc_rows(B, n_sm);
    const long n_tiles = (B + rows - 1) / rows;
    *grid = (int)(n_tiles < n_sm ? n_tiles : n_sm);
    *slots = (long)*grid * rows;
    *rows_out = rows;
    return MATAS_OK;
}

static int decode_use_tab() {
    // MATAS_DECODE_TAB=0: block 0 keeps per-env cache rows instead of the shared tables -- the comparand of the
    // table-mode regression test (scripts/tab_check.py, tests/test_policy_gpu.py)
    static int use_tab = -1;
    if (use_tab < 0) { const char* e = getenv("MATAS_DECODE_TAB"); use_tab = e ? atoi(e) : 1; }
    return use_tab;
}

int launch_decode_tc(const PolicyBufs& b, const DecodeArgs& a, int grid, cudaStream_t st) {
    const size_t bytes = dsmem_bytes();
    // stochastic decode: the standard table; deterministic Semi_Discrete decode: plus the speculative rows of its stride blocks
    const float* tab = (decode_use_tab() && b.x1tab && (!a.deterministic || (a.mode == MATAS_MODE_SEMI && b.x1spec))) ? b.x1tab : nullptr;
    if (tab && a.deterministic) {
        CUDA_TRY(cudaFuncSetAttribute(decode_tc_kernel<PF_ROWS, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
        decode_tc_kernel<PF_ROWS, 2><<<grid, NTD, bytes, st>>>(b, a, tab, b.x1spec);
    } else if (tab) {
        CUDA_TRY(cudaFuncSetAttribute(decode_tc_kernel<PF_ROWS, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
        decode_tc_kernel<PF_ROWS, 1><<<grid, NTD, bytes, st>>>(b, a, tab, nullptr);
    } else {
        CUDA_TRY(cudaFuncSetAttribute(decode_tc_kernel<PF_ROWS, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
        decode_tc_kernel<PF_ROWS, 0><<<grid, NTD, bytes, st>>>(b, a, nullptr, nullptr);
    }
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}
