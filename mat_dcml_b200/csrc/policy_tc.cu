// Tensor-core (tcgen05) token-parallel MAT-AS kernels for sm_100a: 128-token tiles, fp16
// hi + scaled-lo operands in shared memory (tc.cuh), fp32 accumulators in tensor memory, thread-per-row
// epilogues fusing bias / GELU / residual / LayerNorm; the attention of each block is attn_tc.cu's kernel, whose
// output rows are this file's first GEMM operand.  Two CTAs per SM so that one CTA's epilogue overlaps the other's MMAs.
// Token-major [128 x 64] fp32 tiles (residual stream in, every token-major output out) move by tensor-map TMA between
// global memory and a dense swizzled shared tile (tma.cuh, struct TileIO below); weights stream through a two-slot ring of
// 1-D bulk copies fed by the CTA's last thread (thread 0 issues the MMAs).
//
// Reference semantics: EncodeBlock / Encoder of mat_src/mat/algorithms/mat/algorithm/ma_transformer.py:70-154.
#include <stdlib.h>

#include "policy_fwd.h"
#include "tc.cuh"
#include "tma.cuh"

using namespace tile;

namespace {

constexpr int TC_DEPTH = 2;

__constant__ int c_tc_enc_blk0[6] = {lay::hb(lay::enc_blob(0, lay::EP)), lay::hb(lay::enc_blob(0, lay::EM0)),
                                     lay::hb(lay::enc_blob(0, lay::EM2)), lay::hb(lay::enc_blob(1, lay::EQ)),
                                     lay::hb(lay::enc_blob(1, lay::EK)), lay::hb(lay::enc_blob(1, lay::EV))};
__constant__ int c_tc_enc_blk1[6] = {lay::hb(lay::enc_blob(1, lay::EP)), lay::hb(lay::enc_blob(1, lay::EM0)),
                                     lay::hb(lay::enc_blob(1, lay::EM2)), lay::hb(lay::ENC_H0), lay::hb(lay::dec_q2(0)),
                                     lay::hb(lay::dec_q2(1))};
__constant__ int c_tc_enc_blk1_dec[8] = {lay::hb(lay::enc_blob(1, lay::EP)), lay::hb(lay::enc_blob(1, lay::EM0)),
                                         lay::hb(lay::enc_blob(1, lay::EM2)), lay::hb(lay::ENC_H0),
                                         lay::hb(lay::fold_blob(0, lay::FQ0)), lay::hb(lay::fold_blob(0, lay::FQ1)),
                                         lay::hb(lay::fold_blob(1, lay::FQ0)), lay::hb(lay::fold_blob(1, lay::FQ1))};

__constant__ int c_tc_enc_embed[3] = {lay::hb(lay::enc_blob(0, lay::EQ)), lay::hb(lay::enc_blob(0, lay::EK)),
                                      lay::hb(lay::enc_blob(0, lay::EV))};
__constant__ int c_tc_dec_embed[3] = {lay::hb(lay::dec_blob(0, lay::DQ1)), lay::hb(lay::dec_blob(0, lay::DK1)),
                                      lay::hb(lay::dec_blob(0, lay::DV1))};
__constant__ int c_tc_dec_attn1[2][3] = {
    {lay::hb(lay::dec_blob(0, lay::DP1)), lay::hb(lay::dec_blob(0, lay::DK2)), lay::hb(lay::dec_blob(0, lay::DV2))},
    {lay::hb(lay::dec_blob(1, lay::DP1)), lay::hb(lay::dec_blob(1, lay::DK2)), lay::hb(lay::dec_blob(1, lay::DV2))}};
__constant__ int c_tc_dec_attn2[2][6] = {
    {lay::hb(lay::dec_blob(0, lay::DP2)), lay::hb(lay::dec_blob(0, lay::DM0)), lay::hb(lay::dec_blob(0, lay::DM2)),
     lay::hb(lay::dec_blob(1, lay::DQ1)), lay::hb(lay::dec_blob(1, lay::DK1)), lay::hb(lay::dec_blob(1, lay::DV1))},
    {lay::hb(lay::dec_blob(1, lay::DP2)), lay::hb(lay::dec_blob(1, lay::DM0)), lay::hb(lay::dec_blob(1, lay::DM2)),
     lay::hb(lay::DEC_H0), 0, 0}};

__device__ __forceinline__ void add_bias32(float (&v)[32], const float* __restrict__ blob, int half) {
    const float4* bias = reinterpret_cast<const float4*>(blob + lay::B_BIAS + half * 32);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float4 b = bias[j];
        v[j * 4] += b.x; v[j * 4 + 1] += b.y; v[j * 4 + 2] += b.z; v[j * 4 + 3] += b.w;
    }
}

// one GEMM stage: acquire the weights, publish the operand, issue, refill the ring, wait, read the
// accumulator and add the bias.  `blob` stays valid until ring.release().
struct Stage {
    tc::Ctx& c;
    Ring& ring;
    __device__ __forceinline__ const float* run(float (&v)[32]) {
        const float* blob = ring.acquire();
        tc::gemm_issue(c, blob);
        ring.prefetch();
        tc::gemm_result(c, v);
        add_bias32(v, blob, c.half);
        return blob;
    }
};
__device__ __forceinline__ void affine32(float (&v)[32], const float* __restrict__ blob, int half) {
    const float4* w = reinterpret_cast<const float4*>(blob + lay::B_LNW + half * 32);
    const float4* b = reinterpret_cast<const float4*>(blob + lay::B_LNB + half * 32);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float4 ww = w[j], bb = b[j];
        v[j * 4] = fmaf(v[j * 4], ww.x, bb.x); v[j * 4 + 1] = fmaf(v[j * 4 + 1], ww.y, bb.y);
        v[j * 4 + 2] = fmaf(v[j * 4 + 2], ww.z, bb.z); v[j * 4 + 3] = fmaf(v[j * 4 + 3], ww.w, bb.w);
    }
}
__device__ __forceinline__ void save_row32(float* __restrict__ base, long tile0, int row, int half, int n_valid, const float (&v)[32]) {
    if (base && row < n_valid) {                     // thread-per-row: 128 contiguous bytes as four whole-sector stores
        float* p = base + (tile0 + row) * 64 + half * 32;
#pragma unroll
        for (int j = 0; j < 4; ++j)
            st_global_256(p + j * 8, v[j * 8], v[j * 8 + 1], v[j * 8 + 2], v[j * 8 + 3], v[j * 8 + 4], v[j * 8 + 5], v[j * 8 + 6], v[j * 8 + 7]);
    }
}

// -------------------------------------------------------------------------------------------------
// encoder block: attention (bidirectional) + proj + LN1 + MLP + LN2 (ma_transformer.py:89-92);
// BLK 0 then emits QKV of block 1; BLK 1 emits obs_rep, the value head (:138-139) and the decoder's
// cross-attention queries (plain Q2 for the teacher-forced passes, folded Q2' for the decode).
// -------------------------------------------------------------------------------------------------
constexpr int PF_AHEAD = 2 * 148;        // tiles between a CTA and the one that takes its place (2 CTAs on each of 148 SMs)
// L2 prefetch of the 128 token rows (256 B each) of one tile; 256 threads, two 128-byte lines each
__device__ __forceinline__ void prefetch_tile_rows(const float* __restrict__ base, long tile0, long n_tok, int tid) {
    const long g = tile0 + (tid >> 1);
    if (g < n_tok) asm volatile("prefetch.global.L2 [%0];" ::"l"(base + g * 64 + (tid & 1) * 32));
}

// attention output rows computed by attn_tc.cu (token-major, global) -> operand, one half-warp per row
__device__ __forceinline__ void rows_to_operand(unsigned char* sA, const float* __restrict__ y, long tile0, long n_tok) {
    const int hw = threadIdx.x >> 4, l = threadIdx.x & 15;
    constexpr int N = tc::TR * 16 / NT;
    float4 c[N];                                                  // all loads in flight before the first conversion / store
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const long g = tile0 + hw + i * (NT / 16);
        c[i] = g < n_tok ? __ldg(reinterpret_cast<const float4*>(y + g * 64) + l) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int i = 0; i < N; ++i) tc::store_operand4(sA, hw + i * (NT / 16), l, c[i]);
}

// The dense swizzled tile of a CTA (tma.cuh) and the launch's tensor maps: loads and stores of whole token-major tiles by
// one thread (the last: thread 0 issues the MMAs).  A store may be issued after a CTA barrier that follows the writers'
// proxy fence; the tile may be rewritten after free() by the issuing thread and one more CTA barrier.
struct TileIO {
    float* sD;
    uint32_t d0, d1;
    const tma::RowMaps* maps;
    int tile0;
    bool issuer;
    __device__ __forceinline__ void init(float* tile, const tma::RowMaps* m, long t0) {
        sD = tile; d0 = smem_u32(tile); d1 = smem_u32(tile + tma::BOX_FLOATS); maps = m; tile0 = (int)t0; issuer = threadIdx.x == NT - 1;
    }
    __device__ __forceinline__ void load(int mi, uint64_t* bar) const {          // completes on `bar` (32 KB of transaction bytes)
        if (issuer) {
            mbar_expect_tx(bar, 2 * tma::BOX_FLOATS * 4);
            tma::load_box(d0, &maps->m[mi], 0, tile0, smem_u32(bar));
            tma::load_box(d1, &maps->m[mi], 32, tile0, smem_u32(bar));
        }
    }
    __device__ __forceinline__ void store(int mi) const {
        if (issuer) {
            tma::store_box(&maps->m[mi], d0, 0, tile0);
            tma::store_box(&maps->m[mi], d1, 32, tile0);
            tma::store_commit();
        }
    }
    __device__ __forceinline__ void free() const { if (issuer) tma::store_wait_read(); }
    // a stage result -> tile -> global.  The wait for the previous store's reads sits here, a stage's MMA round trip and
    // accumulator read after that store was issued (a 32 KB store takes ~1,500 cycles to leave shared memory), not before
    // the stage's first barrier.
    __device__ __forceinline__ void out(int mi, int row, int half, const float (&v)[32]) const {
        free();
        __syncthreads();
        tma::row_store32(sD, row, half, v);
        tc::fence_proxy_async();
        __syncthreads();
        store(mi);
    }
};

// Token-major tiles move by tensor-map TMA (tma.cuh): the residual stream arrives as two swizzled boxes in the dense tile
// sD while the prologue runs, and every token-major output (x, Q / K / V, obs_rep, the cross-attention queries) leaves from
// sD as two box stores issued by one thread, so no thread spends load / store slots on them and the store drains under the
// next stage's MMAs.  maps.m: [0] X; BLK 0: [1..3] Q, K, V of block 1; BLK 1: [1] obs_rep, [2..] the query outputs.
template <int BLK>
__global__ void __launch_bounds__(NT, 2)
enc_block_tc(PolicyBufs b, TrainBufs tb, float* __restrict__ values, long n_tok, int L, int for_decode, const float* __restrict__ y_in,
             const __grid_constant__ tma::RowMaps maps) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    tc::Smem s = tc::carve(smem_raw, TC_DEPTH);
    const int tid = threadIdx.x;
    const long tile0 = (long)blockIdx.x * tc::TR;
    const int n_valid = (int)min((long)tc::TR, n_tok - tile0);
    // pull this tile's attention rows, and the inputs of the tile that will follow on this SM (CTAs start in index order,
    // two per SM), into L2 while the prologue (TMEM allocation, barrier set-up, first weight blob) runs
    prefetch_tile_rows(y_in, tile0, n_tok, tid);
    prefetch_tile_rows(b.X, tile0 + (long)PF_AHEAD * tc::TR, n_tok, tid);
    prefetch_tile_rows(y_in, tile0 + (long)PF_AHEAD * tc::TR, n_tok, tid);
    Ring ring;
    if (BLK == 1 && for_decode) ring.init(s.ring, s.bars, b.w, c_tc_enc_blk1_dec, 8, 8, TC_DEPTH, lay::BLOB_BYTES, tc::SLOT_F, NT - 1);
    else ring.init(s.ring, s.bars, b.w, BLK == 0 ? c_tc_enc_blk0 : c_tc_enc_blk1, 6, 6, TC_DEPTH, lay::BLOB_BYTES, tc::SLOT_F, NT - 1);
    uint64_t* x_bar = &s.bars[6];
    if (tid == 0) mbar_init(x_bar, 1);
    tc::Ctx c;
    tc::ctx_init(c, s);                  // fences the barrier initialisations; ends with a CTA barrier
    TileIO io;
    io.init(s.sT, &maps, tile0);         // dense swizzled tile: two boxes of [128][32] floats (1024-byte aligned, see tc::carve)
    float* sD = io.sD;
    io.load(0, x_bar);                   // residual stream -> sD
    ring.prefetch();
    if (BLK == 1) for (int i = tid; i < lay::ENC_H3_SIZE; i += NT) s.sSmall[512 + i] = b.w[lay::ENC_H3 + i];
    const int row = c.row, half = c.half;
    float* sRed = s.sSmall;              // [TR][2] float2 LayerNorm partials, [512 + ...) head.3, [640 + ...) head dot partials
    Stage st = {c, ring};
    rows_to_operand(s.sA, y_in, tile0, n_tok);                   // attention output of attn_tc.cu -> first operand
    float v[32], res[32];
    // ---- proj + residual + LN1 ------------------------------------------------------------------
    {
        const float* blob = st.run(v);
        mbar_wait(x_bar, 0);                                      // the residual tile has landed
        tma::row_load32(sD, row, half, res);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] += res[i];
        const float rs = tc::layernorm_pair(v, sRed, row, half);
        if (tb.on) {
            save_row32(tb.e_xh1[BLK], tile0, row, half, n_valid, v);
            if (half == 0 && row < n_valid) tb.e_rs1[BLK][tile0 + row] = rs;
        }
        affine32(v, blob, half);
        ring.release();
        tma::row_store32(sD, row, half, v);                       // x1: residual of the MLP (each thread rewrites what it read)
        tc::store_operand32(s.sA, row, half, v);
    }
    // ---- mlp.0 + GELU ---------------------------------------------------------------------------
    {
        st.run(v);
        ring.release();
        if (tb.on) save_row32(tb.e_m0[BLK], tile0, row, half, n_valid, v);
#pragma unroll
        for (int i = 0; i < 32; i += 2) gelu_erf_x2(v[i], v[i + 1]);
        tc::store_operand32(s.sA, row, half, v);
    }
    // ---- mlp.2 + residual + LN2 -----------------------------------------------------------------
    {
        const float* blob = st.run(v);
        tma::row_load32(sD, row, half, res);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] += res[i];
        const float rs = tc::layernorm_pair(v, sRed, row, half);
        if (tb.on) {
            save_row32(tb.e_xh2[BLK], tile0, row, half, n_valid, v);
            if (half == 0 && row < n_valid) tb.e_rs2[BLK][tile0 + row] = rs;
        }
        affine32(v, blob, half);
        ring.release();
        tma::row_store32(sD, row, half, v);                       // block output
        tc::store_operand32(s.sA, row, half, v);
    }
    if (BLK == 0) {
#pragma unroll
        for (int o = 0; o < 3; ++o) {
            const float* blob = ring.acquire();
            tc::gemm_issue(c, blob);                              // proxy fence + barrier: sD holds the complete x2
            ring.prefetch();
            if (o == 0) io.store(0);                              // x2 -> X
            tc::gemm_result(c, v);
            add_bias32(v, blob, half);
            ring.release();
            io.out(1 + o, row, half, v);
        }
    } else {
        // ---- value head: Linear -> GELU -> LN -> Linear(64 -> 1) --------------------------------------
        {
            const float* blob = ring.acquire();
            tc::gemm_issue(c, blob);
            ring.prefetch();
            io.store(1);                                        // x2 -> obs_rep
            tc::gemm_result(c, v);
            add_bias32(v, blob, half);
            if (tb.on) save_row32(tb.e_hh, tile0, row, half, n_valid, v);
#pragma unroll
            for (int i = 0; i < 32; i += 2) gelu_erf_x2(v[i], v[i + 1]);
            const float rs = tc::layernorm_pair(v, sRed, row, half);
            if (tb.on) {
                save_row32(tb.e_xhh, tile0, row, half, n_valid, v);
                if (half == 0 && row < n_valid) tb.e_rsh[tile0 + row] = rs;
            }
            affine32(v, blob, half);
            ring.release();
            const float* w3 = s.sSmall + 512 + half * 32;
            float d = 0.f;
#pragma unroll
            for (int i = 0; i < 32; ++i) d = fmaf(v[i], w3[i], d);
            float* sDot = s.sSmall + 640;                         // clear of the LayerNorm exchange [0, 512) other rows may still read
            sDot[row * 2 + half] = d;
            __syncthreads();
            if (half == 0 && row < n_valid) values[tile0 + row] = sDot[row * 2] + sDot[row * 2 + 1] + s.sSmall[512 + 64];
        }
        // ---- cross-attention queries from obs_rep (operand still holds it) ------------------------------
        const int n_q = for_decode ? 4 : 2;
        for (int o = 0; o < n_q; ++o) {
            const float* blob = ring.acquire();
            tc::gemm_issue(c, blob);
            ring.prefetch();
            tc::gemm_result(c, v);
            add_bias32(v, blob, half);
            ring.release();
            io.out(2 + o, row, half, v);
        }
    }
    io.free();                                                  // shared memory must outlive the last store's reads
    tc::ctx_fini(c);
}

// Tail shared by the kernels that end with three projections of the tile in the operand (x also in the dense tile):
// x -> maps mx0 / mx1 (token-major; mx1 < 0: none), then out[o] = x . W_o + b_o for the next three ring blobs -> maps mo + o.
__device__ __forceinline__ void qkv_tail(tc::Ctx& c, Ring& ring, const TileIO& io, int mx0, int mx1, int mo) {
    float v[32];
#pragma unroll
    for (int o = 0; o < 3; ++o) {
        const float* blob = ring.acquire();
        tc::gemm_issue(c, blob);                              // proxy fence + barrier: the tile holds the complete x
        ring.prefetch();
        if (o == 0) {
            io.store(mx0);
            if (mx1 >= 0) io.store(mx1);
        }
        tc::gemm_result(c, v);
        add_bias32(v, blob, c.half);
        ring.release();
        io.out(mo + o, c.row, c.half, v);
    }
    io.free();                                                // shared memory must outlive the last store's reads
}

// -------------------------------------------------------------------------------------------------
// encoder stage 1: obs LayerNorm -> Linear(obs_dim->64) -> GELU -> LayerNorm -> X ; QKV of block 0
// (ma_transformer.py:133-136,148-151 and :50-52).  The 7-wide input layer runs on the CUDA cores
// (one thread per row and column half), the three projections on the tensor cores.
// -------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(NT, 2)
enc_embed_qkv_tc(PolicyBufs b, TrainBufs tb, const float* __restrict__ obs, long n_tok, int obs_dim, const __grid_constant__ tma::RowMaps maps) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    tc::Smem s = tc::carve(smem_raw, TC_DEPTH);
    const int tid = threadIdx.x;
    const long tile0 = (long)blockIdx.x * tc::TR;
    const int n_valid = (int)min((long)tc::TR, n_tok - tile0);
    Ring ring;
    ring.init(s.ring, s.bars, b.w, c_tc_enc_embed, 3, 3, TC_DEPTH, lay::BLOB_BYTES, tc::SLOT_F, NT - 1);
    float* sPar = s.sT;                                   // the small parameter group lives in sT until x0 is formed
    for (int i = tid; i < lay::ENC_EMB_SIZE; i += NT) sPar[i] = b.w[lay::ENC_EMB + i];
    // The tile's observations (n_valid rows of obs_dim floats, contiguous in the env-major / agent-minor layout of the
    // policy API) are staged with ONE TMA bulk copy into shared memory behind the parameters instead of eight strided scalar
    // loads per thread; a tile whose byte range is not 16-byte aligned (odd slice offsets, the ragged last tile) keeps the loads.
    float* sObs = s.sT + 1024;
    uint64_t* obs_bar = &s.bars[6];
    const uint32_t obs_bytes = (uint32_t)n_valid * obs_dim * 4u;
    const bool obs_tma = ((reinterpret_cast<uintptr_t>(obs + tile0 * obs_dim) | obs_bytes) & 15) == 0;
    if (tid == 0) mbar_init(obs_bar, 1);
    tc::Ctx c;
    tc::ctx_init(c, s);                                   // fences the barrier initialisations, CTA barrier
    if (tid == 0 && obs_tma) {
        mbar_expect_tx(obs_bar, obs_bytes);
        bulk_g2s(sObs, obs + tile0 * obs_dim, obs_bytes, obs_bar);
    }
    ring.prefetch();
    const int row = c.row, half = c.half;
    float v[32];
    {
        const long g = tile0 + row;
        float o[8], mean = 0.f;
        if (obs_tma) mbar_wait(obs_bar, 0);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            o[k] = (k < obs_dim && row < n_valid) ? (obs_tma ? sObs[row * obs_dim + k] : obs[g * obs_dim + k]) : 0.f;
            mean += o[k];
        }
        mean /= (float)obs_dim;
        float var = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) if (k < obs_dim) var += (o[k] - mean) * (o[k] - mean);
        const float rs7 = rsqrtf(var / (float)obs_dim + 1e-5f);
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = k < obs_dim ? (o[k] - mean) * rs7 * sPar[lay::ENC_EMB_LNW + k] + sPar[lay::ENC_EMB_LNB + k] : 0.f;
        tc::row_load32(sPar + lay::ENC_EMB_B + half * 32, v);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            if (k < obs_dim) {
                float w[32];
                tc::row_load32(sPar + lay::ENC_EMB_WT + k * 64 + half * 32, w);
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = fmaf(o[k], w[i], v[i]);
            }
        }
        if (tb.on) save_row32(tb.e_h0, tile0, row, half, n_valid, v);
#pragma unroll
        for (int i = 0; i < 32; i += 2) gelu_erf_x2(v[i], v[i + 1]);
        const float rs = tc::layernorm_pair(v, s.sSmall, row, half);
        if (tb.on) {
            save_row32(tb.e_xh0, tile0, row, half, n_valid, v);
            if (half == 0 && row < n_valid) tb.e_rs0[tile0 + row] = rs;
        }
        float lw[32];
        tc::row_load32(sPar + lay::ENC_EMB_LN2W + half * 32, lw);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] *= lw[i];
        tc::row_load32(sPar + lay::ENC_EMB_LN2B + half * 32, lw);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] += lw[i];
    }
    __syncthreads();                                       // every thread is done with the parameters in sT
    TileIO io;
    io.init(s.sT, &maps, tile0);
    tma::row_store32(io.sD, row, half, v);
    tc::store_operand32(s.sA, row, half, v);
    qkv_tail(c, ring, io, 0, tb.on ? 1 : -1, 2);          // maps: X, e_x0, Q, K, V of block 0
    tc::ctx_fini(c);
}

// -------------------------------------------------------------------------------------------------
// teacher-forced decoder, stage 1: shifted-action embedding -> X ; QKV1 of block 0
// (transformer_act.py:103-111,176-181)
// -------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(NT, 2)
dec_embed_qkv_tc(PolicyBufs b, TrainBufs tb, const float* __restrict__ actions, long n_tok, int L, const __grid_constant__ tma::RowMaps maps) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    tc::Smem s = tc::carve(smem_raw, TC_DEPTH);
    const long tile0 = (long)blockIdx.x * tc::TR;
    const int n_valid = (int)min((long)tc::TR, n_tok - tile0);
    Ring ring;
    ring.init(s.ring, s.bars, b.w, c_tc_dec_embed, 3, 3, TC_DEPTH, lay::BLOB_BYTES, tc::SLOT_F, NT - 1);
    build_emb_table(s.sSmall + 512, b.w);
    tc::Ctx c;
    tc::ctx_init(c, s);
    ring.prefetch();
    const int row = c.row, half = c.half;
    TileIO io;
    io.init(s.sT, &maps, tile0);
    {
        const long g = tile0 + row;
        int code = 0;
        if (g < n_tok) code = (g % L) == 0 ? 1 : (actions[g - 1] != 0.f ? 3 : 2);
        float v[32];
        tc::row_load32(s.sSmall + 512 + code * 64 + half * 32, v);
        tma::row_store32(io.sD, row, half, v);
        tc::store_operand32(s.sA, row, half, v);
    }
    qkv_tail(c, ring, io, 0, tb.on ? 1 : -1, 2);          // maps: X, d_x0, Q1, K1, V1 of block 0
    tc::ctx_fini(c);
}

// decoder block, first half: causal self-attention + proj + LN1, then K2 / V2 of the result
// (ma_transformer.py:113 and the key/value side of :114)
__global__ void __launch_bounds__(NT, 2)
dec_attn1_tc(PolicyBufs b, TrainBufs tb, int blk, long n_tok, int L, const float* __restrict__ y_in, const __grid_constant__ tma::RowMaps maps) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    tc::Smem s = tc::carve(smem_raw, TC_DEPTH);
    const long tile0 = (long)blockIdx.x * tc::TR;
    const int n_valid = (int)min((long)tc::TR, n_tok - tile0);
    Ring ring;
    ring.init(s.ring, s.bars, b.w, c_tc_dec_attn1[blk], 3, 3, TC_DEPTH, lay::BLOB_BYTES, tc::SLOT_F, NT - 1);
    uint64_t* x_bar = &s.bars[6];
    if (threadIdx.x == 0) mbar_init(x_bar, 1);
    tc::Ctx c;
    tc::ctx_init(c, s);
    TileIO io;
    io.init(s.sT, &maps, tile0);                              // maps: X, K2, V2 of this block
    io.load(0, x_bar);                                        // residual stream
    ring.prefetch();
    const int row = c.row, half = c.half;
    float* sRed = s.sSmall;
    rows_to_operand(s.sA, y_in, tile0, n_tok);
    float v[32], res[32];
    Stage st = {c, ring};
    {
        const float* blob = st.run(v);
        mbar_wait(x_bar, 0);
        tma::row_load32(io.sD, row, half, res);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] += res[i];
        const float rs = tc::layernorm_pair(v, sRed, row, half);
        if (tb.on) {
            save_row32(tb.d_xh1[blk], tile0, row, half, n_valid, v);
            if (half == 0 && row < n_valid) tb.d_rs1[blk][tile0 + row] = rs;
        }
        affine32(v, blob, half);
        ring.release();
        tc::store_operand32(s.sA, row, half, v);                  // x1: input of the cross-attention keys / values
    }
#pragma unroll
    for (int o = 0; o < 2; ++o) {
        const float* blob = ring.acquire();
        tc::gemm_issue(c, blob);                                  // barrier: the residual reads are done
        ring.prefetch();
        tc::gemm_result(c, v);
        add_bias32(v, blob, half);
        ring.release();
        io.out(1 + o, row, half, v);
    }
    io.free();
    tc::ctx_fini(c);
}

// decoder block, second half: cross-attention (query = obs_rep) + proj + LN2(rep + .) + MLP + LN3;
// block 0 then emits QKV1 of block 1, block 1 the logits head and the per-token log-prob / entropy.
__global__ void __launch_bounds__(NT, 2)
dec_attn2_tc(PolicyBufs b, TrainBufs tb, int blk, const float* __restrict__ actions, const float* __restrict__ ava,
             float* __restrict__ logp, float* __restrict__ entropy, long n_tok, int L, int n_disc, const float* __restrict__ y_in,
             const __grid_constant__ tma::RowMaps maps) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    tc::Smem s = tc::carve(smem_raw, TC_DEPTH);
    const int tid = threadIdx.x;
    const long tile0 = (long)blockIdx.x * tc::TR;
    const int n_valid = (int)min((long)tc::TR, n_tok - tile0);
    Ring ring;
    ring.init(s.ring, s.bars, b.w, c_tc_dec_attn2[blk], blk == 0 ? 6 : 4, blk == 0 ? 6 : 4, TC_DEPTH, lay::BLOB_BYTES, tc::SLOT_F, NT - 1);
    uint64_t* x_bar = &s.bars[6];
    if (tid == 0) mbar_init(x_bar, 1);
    tc::Ctx c;
    tc::ctx_init(c, s);
    TileIO io;
    io.init(s.sT, &maps, tile0);                              // maps: obs_rep; block 0: X, Q1 / K1 / V1 of block 1
    io.load(0, x_bar);                                        // obs_rep: the residual of LN2
    ring.prefetch();
    if (blk == 1) for (int i = tid; i < lay::DEC_H3_SIZE; i += NT) s.sSmall[512 + i] = b.w[lay::DEC_H3 + i];
    const int row = c.row, half = c.half;
    float* sRed = s.sSmall;
    rows_to_operand(s.sA, y_in, tile0, n_tok);
    float v[32], res[32];
    Stage st = {c, ring};
    {   // proj + obs_rep + LN2
        const float* blob = st.run(v);
        mbar_wait(x_bar, 0);
        tma::row_load32(io.sD, row, half, res);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] += res[i];
        const float rs = tc::layernorm_pair(v, sRed, row, half);
        if (tb.on) {
            save_row32(tb.d_xh2[blk], tile0, row, half, n_valid, v);
            if (half == 0 && row < n_valid) tb.d_rs2[blk][tile0 + row] = rs;
        }
        affine32(v, blob, half);
        ring.release();
        tma::row_store32(io.sD, row, half, v);
        tc::store_operand32(s.sA, row, half, v);
    }
    {   // mlp.0 + GELU
        st.run(v);
        ring.release();
        if (tb.on) save_row32(tb.d_m0[blk], tile0, row, half, n_valid, v);
#pragma unroll
        for (int i = 0; i < 32; i += 2) gelu_erf_x2(v[i], v[i + 1]);
        tc::store_operand32(s.sA, row, half, v);
    }
    {   // mlp.2 + residual + LN3
        const float* blob = st.run(v);
        tma::row_load32(io.sD, row, half, res);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] += res[i];
        const float rs = tc::layernorm_pair(v, sRed, row, half);
        if (tb.on) {
            save_row32(tb.d_xh3[blk], tile0, row, half, n_valid, v);
            if (half == 0 && row < n_valid) tb.d_rs3[blk][tile0 + row] = rs;
        }
        affine32(v, blob, half);
        ring.release();
        tma::row_store32(io.sD, row, half, v);
        tc::store_operand32(s.sA, row, half, v);
    }
    if (blk == 0) {
        qkv_tail(c, ring, io, 1, -1, 2);
    } else {
        // ---- logits head + per-token log-prob / entropy (transformer_act.py:112-128) -----------------------
        const float* blob = st.run(v);
        if (tb.on) save_row32(tb.d_hh, tile0, row, half, n_valid, v);
#pragma unroll
        for (int i = 0; i < 32; i += 2) gelu_erf_x2(v[i], v[i + 1]);
        const float rs = tc::layernorm_pair(v, sRed, row, half);
        if (tb.on) {
            save_row32(tb.d_xhh, tile0, row, half, n_valid, v);
            if (half == 0 && row < n_valid) tb.d_rsh[tile0 + row] = rs;
        }
        affine32(v, blob, half);
        ring.release();
        const float* w3 = s.sSmall + 512;
        float d0 = 0.f, d1 = 0.f;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            d0 = fmaf(v[i], w3[half * 32 + i], d0);
            d1 = fmaf(v[i], w3[64 + half * 32 + i], d1);
        }
        float* sLog = s.sT;                                        // [128][2 halves][2 logits]; the x3 tile is dead in block 1
        sLog[(row * 2 + half) * 2] = d0;
        sLog[(row * 2 + half) * 2 + 1] = d1;
        __syncthreads();
        if (half == 0 && row < n_valid) {
            const long g = tile0 + row;
            const int pos = (int)(g % L);
            const float l0 = sLog[row * 4] + sLog[row * 4 + 2] + w3[lay::DEC_H3_B];
            const float l1 = sLog[row * 4 + 1] + sLog[row * 4 + 3] + w3[lay::DEC_H3_B + 1];
            const float a = actions[g];
            if (pos < n_disc) {
                float z0, z1;
                masked_logz(l0, l1, ava ? ava + g * 2 : nullptr, z0, z1);
                logp[g] = a != 0.f ? z1 : z0;
                const float c0 = fmaxf(z0, -3.4028234663852886e38f), c1 = fmaxf(z1, -3.4028234663852886e38f);
                entropy[g] = -(expf(c0) * c0 + expf(c1) * c1);
            } else {
                const float sd = 0.5f / (1.0f + expf(-w3[lay::DEC_H3_LOGSTD + 1]));
                const float d = a - l1;
                logp[g] = -(d * d) / (2.0f * sd * sd) - logf(sd) - HALF_LOG_2PI;
                entropy[g] = 0.5f + HALF_LOG_2PI + logf(sd);
            }
        }
    }
    tc::ctx_fini(c);
}

}  // namespace

// tensor maps of up to MAX_MAPS token-major arrays of this launch (null entries are never used by the kernel)
static int build_maps(tma::RowMaps& maps, const float* const (&arrs)[tma::MAX_MAPS], long n_tok) {
    for (int i = 0; i < tma::MAX_MAPS; ++i) {
        const int rc = tma::rows_map(&maps.m[i], arrs[i], n_tok);
        if (rc) return rc;
    }
    return MATAS_OK;
}

int launch_enc_embed_tc(const PolicyBufs& b, const TrainBufs& tb, const float* obs, long n_tok, int obs_dim, cudaStream_t st) {
    const size_t bytes = tc::smem_bytes(TC_DEPTH);
    CUDA_TRY(cudaFuncSetAttribute(enc_embed_qkv_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    tma::RowMaps maps;
    const float* const arrs[tma::MAX_MAPS] = {b.X, tb.on ? tb.e_x0 : nullptr, tb.on ? tb.e_q[0] : b.Q, b.K[0], b.V[0], nullptr};
    const int rc = build_maps(maps, arrs, n_tok);
    if (rc) return rc;
    enc_embed_qkv_tc<<<(unsigned)((n_tok + tc::TR - 1) / tc::TR), NT, bytes, st>>>(b, tb, obs, n_tok, obs_dim, maps);
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}

int launch_enc_blocks_tc(const PolicyBufs& b, const TrainBufs& tb, float* values, long n_tok, int L, int for_decode, int n_sm,
                         cudaStream_t st) {
    const size_t bytes = tc::smem_bytes(TC_DEPTH);
    CUDA_TRY(cudaFuncSetAttribute(enc_block_tc<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    CUDA_TRY(cudaFuncSetAttribute(enc_block_tc<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    const unsigned grid = (unsigned)((n_tok + tc::TR - 1) / tc::TR);
    int rc;
    for (int blk = 0; blk < 2; ++blk) {
        float* y = tb.on ? tb.e_y[blk] : b.Y;
        if ((rc = launch_attn_tc(tb.on ? tb.e_q[blk] : b.Q, b.K[blk], b.V[blk], y, tb.on ? tb.e_lse[blk] : nullptr, n_tok / L, L, 0, n_sm, st)))
            return rc;
        tma::RowMaps maps;
        const float* const blk0[tma::MAX_MAPS] = {b.X, tb.on ? tb.e_q[1] : b.Q, b.K[1], b.V[1], nullptr, nullptr};
        const float* const blk1d[tma::MAX_MAPS] = {b.X, b.rep, b.Q2p[0][0], b.Q2p[0][1], b.Q2p[1][0], b.Q2p[1][1]};
        const float* const blk1[tma::MAX_MAPS] = {b.X, b.rep, b.Q2[0], b.Q2[1], nullptr, nullptr};
        if ((rc = build_maps(maps, blk == 0 ? blk0 : (for_decode ? blk1d : blk1), n_tok))) return rc;
        if (blk == 0) enc_block_tc<0><<<grid, NT, bytes, st>>>(b, tb, values, n_tok, L, 0, y, maps);
        else enc_block_tc<1><<<grid, NT, bytes, st>>>(b, tb, values, n_tok, L, for_decode, y, maps);
    }
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}

int launch_teacher_forced_tc(const PolicyBufs& b, const TrainBufs& tb, const float* actions, const float* ava, float* logp,
                             float* entropy, long n_tok, int L, int n_disc, int n_sm, cudaStream_t st) {
    const size_t bytes = tc::smem_bytes(TC_DEPTH);
    CUDA_TRY(cudaFuncSetAttribute(dec_embed_qkv_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    CUDA_TRY(cudaFuncSetAttribute(dec_attn1_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    CUDA_TRY(cudaFuncSetAttribute(dec_attn2_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    const unsigned grid = (unsigned)((n_tok + tc::TR - 1) / tc::TR);
    const long n_env = n_tok / L;
    int rc;
    tma::RowMaps maps;
    {
        const float* const arrs[tma::MAX_MAPS] = {b.X, tb.on ? tb.d_x0 : nullptr, tb.on ? tb.d_q1[0] : b.Q, b.K1[0], b.V1[0], nullptr};
        if ((rc = build_maps(maps, arrs, n_tok))) return rc;
    }
    dec_embed_qkv_tc<<<grid, NT, bytes, st>>>(b, tb, actions, n_tok, L, maps);
    for (int blk = 0; blk < 2; ++blk) {
        float* y = tb.on ? tb.d_y1[blk] : b.Y;
        if ((rc = launch_attn_tc(tb.on ? tb.d_q1[blk] : b.Q, b.K1[blk], b.V1[blk], y, tb.on ? tb.d_lse1[blk] : nullptr, n_env, L, 1, n_sm, st)))
            return rc;
        {
            const float* const arrs[tma::MAX_MAPS] = {b.X, b.K2[blk], b.V2[blk], nullptr, nullptr, nullptr};
            if ((rc = build_maps(maps, arrs, n_tok))) return rc;
        }
        dec_attn1_tc<<<grid, NT, bytes, st>>>(b, tb, blk, n_tok, L, y, maps);
        y = tb.on ? tb.d_y2[blk] : b.Y;
        if ((rc = launch_attn_tc(b.Q2[blk], b.K2[blk], b.V2[blk], y, tb.on ? tb.d_lse2[blk] : nullptr, n_env, L, 1, n_sm, st))) return rc;
        {
            const float* const arrs[tma::MAX_MAPS] = {b.rep, b.X, tb.on ? tb.d_q1[1] : b.Q, b.K1[1], b.V1[1], nullptr};
            if ((rc = build_maps(maps, arrs, n_tok))) return rc;
        }
        dec_attn2_tc<<<grid, NT, bytes, st>>>(b, tb, blk, actions, ava, logp, entropy, n_tok, L, n_disc, y, maps);
    }
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}
