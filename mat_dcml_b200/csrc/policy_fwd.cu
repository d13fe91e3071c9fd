// MAT-AS policy forward kernels for sm_100a (fp32 SIMT GEMMs fed by a TMA bulk-copy weight ring).
//
//   encoder      : enc_embed_qkv -> enc_block<0> -> enc_block<1>            (token-parallel tiles)
//   evaluate     : encoder + dec_embed_qkv -> {dec_attn1 -> dec_attn2} x 2   (teacher forcing)
//   get_actions  : encoder + decode_kernel (persistent, 64 envs per CTA, sequential over agents,
//                  K/V cache in global scratch, stride-block deterministic schedule in-kernel)
//
// Reference semantics: mat_src/mat/algorithms/mat/algorithm/ma_transformer.py:24-230 and
// mat_src/mat/algorithms/utils/transformer_act.py:8-189.
#include <stdlib.h>
#include <string.h>

#include "policy_fwd.h"
#include "tile.cuh"

using namespace tile;

namespace {


// dynamic shared memory carve-up common to all tile kernels
struct Smem {
    float* ring;       // depth * BLOB
    float* sX;         // stream
    float* sA;
    float* sB;
    float* sSmall;     // 1280 floats of small params / scratch
    const float** kptr;
    const float** vptr;
    int* nt;
    uint32_t* bits;    // TM*8 words: per-env action bitset (decode)
    uint64_t* bars;    // ring barriers
};

__host__ __device__ inline size_t smem_bytes(int tm, int depth) {
    size_t f = (size_t)depth * lay::BLOB + 3 * tm * XS + 1280;
    f = (f + 3) & ~(size_t)3;
    return f * 4 + tm * 8 * 2 + tm * 4 + tm * 8 * 4 + 8 * 8 + 16;
}

template <int RPT>
__device__ inline Smem carve(unsigned char* base, int depth) {
    constexpr int TM = 16 * RPT;
    constexpr int TILE_F = TM * XS;
    Smem s;
    float* f = reinterpret_cast<float*>(base);
    s.ring = f; f += (size_t)depth * lay::BLOB;
    s.sX = f; f += TILE_F;
    s.sA = f; f += TILE_F;
    s.sB = f; f += TILE_F;
    s.sSmall = f; f += 1280;
    uintptr_t p = (reinterpret_cast<uintptr_t>(f) + 15) & ~(uintptr_t)15;
    s.bars = reinterpret_cast<uint64_t*>(p); p += 8 * 8;
    s.kptr = reinterpret_cast<const float**>(p); p += TM * 8;
    s.vptr = reinterpret_cast<const float**>(p); p += TM * 8;
    s.nt = reinterpret_cast<int*>(p); p += TM * 4;
    s.bits = reinterpret_cast<uint32_t*>(p);
    return s;
}

template <int RPT>
__device__ __forceinline__ void store_tile(float* __restrict__ dst, long stride, const float* __restrict__ src, int n_valid) {
    for (int idx = threadIdx.x; idx < 16 * RPT * 16; idx += NT) {
        const int r = idx >> 4, c = idx & 15;
        if (r < n_valid)
            *reinterpret_cast<float4*>(dst + (long)r * stride + c * 4) = *reinterpret_cast<const float4*>(src + r * XS + c * 4);
    }
}

// blob sequences (float offsets into the packed weights)
__constant__ int c_seq_enc_embed[3] = {lay::enc_blob(0, lay::EQ), lay::enc_blob(0, lay::EK), lay::enc_blob(0, lay::EV)};
__constant__ int c_seq_enc_blk0[6] = {lay::enc_blob(0, lay::EP), lay::enc_blob(0, lay::EM0), lay::enc_blob(0, lay::EM2),
                                      lay::enc_blob(1, lay::EQ), lay::enc_blob(1, lay::EK), lay::enc_blob(1, lay::EV)};
__constant__ int c_seq_enc_blk1[6] = {lay::enc_blob(1, lay::EP), lay::enc_blob(1, lay::EM0), lay::enc_blob(1, lay::EM2),
                                      lay::ENC_H0, lay::dec_q2(0), lay::dec_q2(1)};
__constant__ int c_seq_dec_embed[3] = {lay::dec_blob(0, lay::DQ1), lay::dec_blob(0, lay::DK1), lay::dec_blob(0, lay::DV1)};
__constant__ int c_seq_dec_attn1[2][3] = {
    {lay::dec_blob(0, lay::DP1), lay::dec_blob(0, lay::DK2), lay::dec_blob(0, lay::DV2)},
    {lay::dec_blob(1, lay::DP1), lay::dec_blob(1, lay::DK2), lay::dec_blob(1, lay::DV2)}};
__constant__ int c_seq_dec_attn2[2][6] = {
    {lay::dec_blob(0, lay::DP2), lay::dec_blob(0, lay::DM0), lay::dec_blob(0, lay::DM2), lay::dec_blob(1, lay::DQ1),
     lay::dec_blob(1, lay::DK1), lay::dec_blob(1, lay::DV1)},
    {lay::dec_blob(1, lay::DP2), lay::dec_blob(1, lay::DM0), lay::dec_blob(1, lay::DM2), lay::DEC_H0, 0, 0}};
__constant__ int c_seq_q2[2] = {lay::dec_q2(0), lay::dec_q2(1)};
// folded cross-attention query maps for the decode: q2'_h of block 0, then block 1
__constant__ int c_seq_q2p[4] = {lay::fold_blob(0, lay::FQ0), lay::fold_blob(0, lay::FQ1), lay::fold_blob(1, lay::FQ0),
                                 lay::fold_blob(1, lay::FQ1)};
__constant__ int c_seq_enc_blk1_dec[8] = {lay::enc_blob(1, lay::EP), lay::enc_blob(1, lay::EM0), lay::enc_blob(1, lay::EM2),
                                          lay::ENC_H0, lay::fold_blob(0, lay::FQ0), lay::fold_blob(0, lay::FQ1),
                                          lay::fold_blob(1, lay::FQ0), lay::fold_blob(1, lay::FQ1)};
// decode step: per block FA0 FA1 FB0 FB1 FC0 FC1 M0 M2, then head.0
__constant__ int c_seq_decode[lay::DEC_FOLD_STEP_BLOBS] = {
    lay::fold_blob(0, lay::FA0), lay::fold_blob(0, lay::FA1), lay::fold_blob(0, lay::FB0), lay::fold_blob(0, lay::FB1),
    lay::fold_blob(0, lay::FC0), lay::fold_blob(0, lay::FC1), lay::dec_blob(0, lay::DM0), lay::dec_blob(0, lay::DM2),
    lay::fold_blob(1, lay::FA0), lay::fold_blob(1, lay::FA1), lay::fold_blob(1, lay::FB0), lay::fold_blob(1, lay::FB1),
    lay::fold_blob(1, lay::FC0), lay::fold_blob(1, lay::FC1), lay::dec_blob(1, lay::DM0), lay::dec_blob(1, lay::DM2),
    lay::DEC_H0};

// per-row metadata for token-parallel tiles: row r <-> token g = tile0 + r = env*L + pos
template <int RPT>
__device__ __forceinline__ void token_rows(const Smem& s, long tile0, long n_tok, int L, const float* K, const float* V,
                                           bool causal) {
    const int r = threadIdx.x;
    if (r < 16 * RPT) {
        const long g = tile0 + r;
        if (g < n_tok) {
            const long env = g / L;
            const int pos = (int)(g - env * L);
            s.kptr[r] = K + env * L * 64;
            s.vptr[r] = V + env * L * 64;
            s.nt[r] = causal ? pos + 1 : L;
        } else {
            s.kptr[r] = K; s.vptr[r] = V; s.nt[r] = 0;
        }
    }
}

// -------------------------------------------------------------------------------------------------
// encoder stage 1: obs LayerNorm -> Linear(obs_dim->64) -> GELU -> LayerNorm -> X ; QKV of block 0
// (ma_transformer.py:133-136,148-151 and :50-52)
// -------------------------------------------------------------------------------------------------
template <int RPT>
__global__ void __launch_bounds__(NT, 2)
enc_embed_qkv(PolicyBufs b, TrainBufs tb, const float* __restrict__ obs, long n_tok, int obs_dim, int depth) {
    constexpr int TM = 16 * RPT;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem s = carve<RPT>(smem_raw, depth);
    const int tid = threadIdx.x, rg = tid >> 4, cg = tid & 15;
    const long tile0 = (long)blockIdx.x * TM;
    const int n_valid = (int)min((long)TM, n_tok - tile0);
    Ring ring;
    ring.init(s.ring, s.bars, b.w, c_seq_enc_embed, 3, 3, depth);
    for (int i = tid; i < lay::ENC_EMB_SIZE; i += NT) s.sSmall[i] = b.w[lay::ENC_EMB + i];
    __syncthreads();
    ring.prefetch();
    float* sObs = s.sSmall + 768;   // [TM][8]
    if (tid < TM) {
        float v[8];
        float mean = 0.f;
        for (int k = 0; k < 8; ++k) {
            v[k] = (k < obs_dim && tid < n_valid) ? obs[(tile0 + tid) * obs_dim + k] : 0.f;
            mean += v[k];
        }
        mean /= (float)obs_dim;
        float var = 0.f;
        for (int k = 0; k < obs_dim; ++k) var += (v[k] - mean) * (v[k] - mean);
        const float rs = rsqrtf(var / (float)obs_dim + 1e-5f);
        for (int k = 0; k < 8; ++k)
            sObs[tid * 8 + k] = k < obs_dim ? (v[k] - mean) * rs * s.sSmall[lay::ENC_EMB_LNW + k] + s.sSmall[lay::ENC_EMB_LNB + k] : 0.f;
    }
    __syncthreads();
    {
        const float* Wt = s.sSmall + lay::ENC_EMB_WT;
        const float4 bb = *reinterpret_cast<const float4*>(s.sSmall + lay::ENC_EMB_B + cg * 4);
        const float4 lw = *reinterpret_cast<const float4*>(s.sSmall + lay::ENC_EMB_LN2W + cg * 4);
        const float4 lb = *reinterpret_cast<const float4*>(s.sSmall + lay::ENC_EMB_LN2B + cg * 4);
        for (int i = 0; i < RPT; ++i) {
            const int r = rg * RPT + i;
            float v0 = bb.x, v1 = bb.y, v2 = bb.z, v3 = bb.w;
            for (int k = 0; k < obs_dim; ++k) {
                const float o = sObs[r * 8 + k];
                const float4 w = *reinterpret_cast<const float4*>(Wt + k * 64 + cg * 4);
                v0 = fmaf(o, w.x, v0); v1 = fmaf(o, w.y, v1); v2 = fmaf(o, w.z, v2); v3 = fmaf(o, w.w, v3);
            }
            if (tb.on && r < n_valid) *reinterpret_cast<float4*>(tb.e_h0 + (tile0 + r) * 64 + cg * 4) = make_float4(v0, v1, v2, v3);
            v0 = gelu_erf(v0); v1 = gelu_erf(v1); v2 = gelu_erf(v2); v3 = gelu_erf(v3);
            const float mean = half_warp_sum(v0 + v1 + v2 + v3) * (1.0f / 64.0f);
            const float d0 = v0 - mean, d1 = v1 - mean, d2 = v2 - mean, d3 = v3 - mean;
            const float var = half_warp_sum(d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3) * (1.0f / 64.0f);
            const float rs = rsqrtf(var + 1e-5f);
            if (tb.on && r < n_valid) {
                *reinterpret_cast<float4*>(tb.e_xh0 + (tile0 + r) * 64 + cg * 4) = make_float4(d0 * rs, d1 * rs, d2 * rs, d3 * rs);
                if (cg == 0) tb.e_rs0[tile0 + r] = rs;
            }
            const float4 o4 = make_float4(d0 * rs * lw.x + lb.x, d1 * rs * lw.y + lb.y, d2 * rs * lw.z + lb.z, d3 * rs * lw.w + lb.w);
            *reinterpret_cast<float4*>(s.sX + r * XS + cg * 4) = o4;
            if (r < n_valid) *reinterpret_cast<float4*>(b.X + (tile0 + r) * 64 + cg * 4) = o4;
            if (tb.on && r < n_valid) *reinterpret_cast<float4*>(tb.e_x0 + (tile0 + r) * 64 + cg * 4) = o4;
        }
    }
    __syncthreads();
    stage<0, RPT>(ring, s.sX, (tb.on ? tb.e_q[0] : b.Q) + tile0 * 64, 64, n_valid, rg, cg);
    stage<0, RPT>(ring, s.sX, b.K[0] + tile0 * 64, 64, n_valid, rg, cg);
    stage<0, RPT>(ring, s.sX, b.V[0] + tile0 * 64, 64, n_valid, rg, cg);
}

// -------------------------------------------------------------------------------------------------
// encoder block: attention (bidirectional) + proj + LN1 + MLP + LN2 (ma_transformer.py:89-92);
// BLK 0 then emits QKV of block 1; BLK 1 emits obs_rep, the value head (:138-139) and the decoder's
// cross-attention queries Q2 = attn2.query(obs_rep) for both decoder blocks (:114, :51).
// -------------------------------------------------------------------------------------------------
template <int BLK, int RPT>
__global__ void __launch_bounds__(NT, 2)
enc_block(PolicyBufs b, TrainBufs tb, float* __restrict__ values, long n_tok, int L, int depth, int for_decode) {
    constexpr int TM = 16 * RPT;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem s = carve<RPT>(smem_raw, depth);
    const int tid = threadIdx.x, rg = tid >> 4, cg = tid & 15;
    const long tile0 = (long)blockIdx.x * TM;
    const int n_valid = (int)min((long)TM, n_tok - tile0);
    Ring ring;
    if (BLK == 1 && for_decode) ring.init(s.ring, s.bars, b.w, c_seq_enc_blk1_dec, 8, 8, depth);
    else ring.init(s.ring, s.bars, b.w, BLK == 0 ? c_seq_enc_blk0 : c_seq_enc_blk1, 6, 6, depth);
    token_rows<RPT>(s, tile0, n_tok, L, b.K[BLK], b.V[BLK], false);
    load_tile<RPT>(s.sX, b.X + tile0 * 64, 64, n_valid);
    load_tile<RPT>(s.sA, (tb.on ? tb.e_q[BLK] : b.Q) + tile0 * 64, 64, n_valid);
    if (BLK == 1) for (int i = tid; i < lay::ENC_H3_SIZE; i += NT) s.sSmall[i] = b.w[lay::ENC_H3 + i];
    __syncthreads();
    ring.prefetch();
    attention_rows<RPT, true>(s.sA, s.sB, s.kptr, s.vptr, s.nt, tb.on ? tb.e_lse[BLK] + tile0 * 2 : nullptr, n_valid);
    SaveOut sv1 = {nullptr, nullptr, nullptr}, sv2 = sv1, sv3 = sv1, sv4 = sv1;
    if (tb.on) {
        store_tile<RPT>(tb.e_y[BLK] + tile0 * 64, 64, s.sB, n_valid);
        sv1.xhat = tb.e_xh1[BLK] + tile0 * 64; sv1.rstd = tb.e_rs1[BLK] + tile0;
        sv2.pre = tb.e_m0[BLK] + tile0 * 64;
        sv3.xhat = tb.e_xh2[BLK] + tile0 * 64; sv3.rstd = tb.e_rs2[BLK] + tile0;
        sv4.pre = tb.e_hh + tile0 * 64; sv4.xhat = tb.e_xhh + tile0 * 64; sv4.rstd = tb.e_rsh + tile0;
    }
    stage<EP_RES | EP_LN, RPT>(ring, s.sB, s.sX, XS, TM, s.sX, XS, TM, rg, cg, &sv1, n_valid);
    __syncthreads(); ring.prefetch();
    stage<EP_GELU, RPT>(ring, s.sX, s.sA, XS, TM, rg, cg, &sv2, n_valid);
    __syncthreads(); ring.prefetch();
    stage<EP_RES | EP_LN, RPT>(ring, s.sA, s.sX, XS, TM, s.sX, XS, TM, rg, cg, &sv3, n_valid);
    __syncthreads(); ring.prefetch();
    if (BLK == 0) {
        store_tile<RPT>(b.X + tile0 * 64, 64, s.sX, n_valid);
        stage<0, RPT>(ring, s.sX, (tb.on ? tb.e_q[1] : b.Q) + tile0 * 64, 64, n_valid, rg, cg);
        stage<0, RPT>(ring, s.sX, b.K[1] + tile0 * 64, 64, n_valid, rg, cg);
        stage<0, RPT>(ring, s.sX, b.V[1] + tile0 * 64, 64, n_valid, rg, cg);
    } else {
        store_tile<RPT>(b.rep + tile0 * 64, 64, s.sX, n_valid);
        stage<EP_GELU | EP_LN, RPT>(ring, s.sX, s.sA, XS, TM, rg, cg, &sv4, n_valid);
        if (for_decode) {
            stage<0, RPT>(ring, s.sX, b.Q2p[0][0] + tile0 * 64, 64, n_valid, rg, cg);
            stage<0, RPT>(ring, s.sX, b.Q2p[0][1] + tile0 * 64, 64, n_valid, rg, cg);
            __syncthreads(); ring.prefetch();
            stage<0, RPT>(ring, s.sX, b.Q2p[1][0] + tile0 * 64, 64, n_valid, rg, cg);
            stage<0, RPT>(ring, s.sX, b.Q2p[1][1] + tile0 * 64, 64, n_valid, rg, cg);
        } else {
            stage<0, RPT>(ring, s.sX, b.Q2[0] + tile0 * 64, 64, n_valid, rg, cg);
            stage<0, RPT>(ring, s.sX, b.Q2[1] + tile0 * 64, 64, n_valid, rg, cg);
        }
        __syncthreads();
        if (tid < n_valid) values[tile0 + tid] = row_dot(s.sA, tid, s.sSmall) + s.sSmall[64];
    }
}

// cross-attention queries from a caller-supplied obs_rep (the transformer_act entry points start
// from obs_rep, transformer_act.py:30,103,131,176): rep_in -> b.rep copy, Q2[blk] = attn2.query(rep)
template <int RPT>
__global__ void __launch_bounds__(NT, 2)
q2_from_rep(PolicyBufs b, const float* __restrict__ rep_in, long n_tok, int depth, int for_decode) {
    constexpr int TM = 16 * RPT;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem s = carve<RPT>(smem_raw, depth);
    const int tid = threadIdx.x, rg = tid >> 4, cg = tid & 15;
    const long tile0 = (long)blockIdx.x * TM;
    const int n_valid = (int)min((long)TM, n_tok - tile0);
    Ring ring;
    if (for_decode) ring.init(s.ring, s.bars, b.w, c_seq_q2p, 4, 4, depth);
    else ring.init(s.ring, s.bars, b.w, c_seq_q2, 2, 2, depth);
    load_tile<RPT>(s.sX, rep_in + tile0 * 64, 64, n_valid);
    __syncthreads();
    ring.prefetch();
    if (rep_in != b.rep) store_tile<RPT>(b.rep + tile0 * 64, 64, s.sX, n_valid);
    if (for_decode) {
        stage<0, RPT>(ring, s.sX, b.Q2p[0][0] + tile0 * 64, 64, n_valid, rg, cg);
        stage<0, RPT>(ring, s.sX, b.Q2p[0][1] + tile0 * 64, 64, n_valid, rg, cg);
        __syncthreads(); ring.prefetch();
        stage<0, RPT>(ring, s.sX, b.Q2p[1][0] + tile0 * 64, 64, n_valid, rg, cg);
        stage<0, RPT>(ring, s.sX, b.Q2p[1][1] + tile0 * 64, 64, n_valid, rg, cg);
    } else {
        stage<0, RPT>(ring, s.sX, b.Q2[0] + tile0 * 64, 64, n_valid, rg, cg);
        stage<0, RPT>(ring, s.sX, b.Q2[1] + tile0 * 64, 64, n_valid, rg, cg);
    }
}

// -------------------------------------------------------------------------------------------------
// teacher-forced decoder, stage 1: shifted-action embedding -> X ; QKV1 of block 0
// (transformer_act.py:103-111,176-181)
// -------------------------------------------------------------------------------------------------
template <int RPT>
__global__ void __launch_bounds__(NT, 2)
dec_embed_qkv(PolicyBufs b, TrainBufs tb, const float* __restrict__ actions, long n_tok, int L, int depth) {
    constexpr int TM = 16 * RPT;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem s = carve<RPT>(smem_raw, depth);
    const int tid = threadIdx.x, rg = tid >> 4, cg = tid & 15;
    const long tile0 = (long)blockIdx.x * TM;
    const int n_valid = (int)min((long)TM, n_tok - tile0);
    Ring ring;
    ring.init(s.ring, s.bars, b.w, c_seq_dec_embed, 3, 3, depth);
    build_emb_table(s.sSmall, b.w);
    __syncthreads();
    ring.prefetch();
    for (int idx = tid; idx < TM * 16; idx += NT) {
        const int r = idx >> 4, c = idx & 15;
        const long g = tile0 + r;
        int code = 0;
        if (g < n_tok) {
            const int pos = (int)(g % L);
            code = pos == 0 ? 1 : (actions[g - 1] != 0.f ? 3 : 2);
        }
        const float4 v = *reinterpret_cast<const float4*>(s.sSmall + code * 64 + c * 4);
        *reinterpret_cast<float4*>(s.sX + r * XS + c * 4) = v;
        if (r < n_valid) *reinterpret_cast<float4*>(b.X + g * 64 + c * 4) = v;
        if (tb.on && r < n_valid) *reinterpret_cast<float4*>(tb.d_x0 + g * 64 + c * 4) = v;
    }
    __syncthreads();
    stage<0, RPT>(ring, s.sX, (tb.on ? tb.d_q1[0] : b.Q) + tile0 * 64, 64, n_valid, rg, cg);
    stage<0, RPT>(ring, s.sX, b.K1[0] + tile0 * 64, 64, n_valid, rg, cg);
    stage<0, RPT>(ring, s.sX, b.V1[0] + tile0 * 64, 64, n_valid, rg, cg);
}

// decoder block, first half: causal self-attention + proj + LN1, then K2/V2 of the result
// (ma_transformer.py:113 and the key/value side of :114)
template <int RPT>
__global__ void __launch_bounds__(NT, 2)
dec_attn1(PolicyBufs b, TrainBufs tb, int blk, long n_tok, int L, int depth) {
    constexpr int TM = 16 * RPT;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem s = carve<RPT>(smem_raw, depth);
    const int tid = threadIdx.x, rg = tid >> 4, cg = tid & 15;
    const long tile0 = (long)blockIdx.x * TM;
    const int n_valid = (int)min((long)TM, n_tok - tile0);
    Ring ring;
    ring.init(s.ring, s.bars, b.w, c_seq_dec_attn1[blk], 3, 3, depth);
    token_rows<RPT>(s, tile0, n_tok, L, b.K1[blk], b.V1[blk], true);
    load_tile<RPT>(s.sX, b.X + tile0 * 64, 64, n_valid);
    load_tile<RPT>(s.sA, (tb.on ? tb.d_q1[blk] : b.Q) + tile0 * 64, 64, n_valid);
    __syncthreads();
    ring.prefetch();
    attention_rows<RPT, true>(s.sA, s.sB, s.kptr, s.vptr, s.nt, tb.on ? tb.d_lse1[blk] + tile0 * 2 : nullptr, n_valid);
    SaveOut sv1 = {nullptr, nullptr, nullptr};
    if (tb.on) {
        store_tile<RPT>(tb.d_y1[blk] + tile0 * 64, 64, s.sB, n_valid);
        sv1.xhat = tb.d_xh1[blk] + tile0 * 64; sv1.rstd = tb.d_rs1[blk] + tile0;
    }
    stage<EP_RES | EP_LN, RPT>(ring, s.sB, s.sX, XS, TM, s.sX, XS, TM, rg, cg, &sv1, n_valid);
    __syncthreads();
    stage<0, RPT>(ring, s.sX, b.K2[blk] + tile0 * 64, 64, n_valid, rg, cg);
    stage<0, RPT>(ring, s.sX, b.V2[blk] + tile0 * 64, 64, n_valid, rg, cg);
}

// per-token distribution head shared by evaluate and decode (transformer_act.py:8-28,112-128)
struct HeadParams { float w0[64]; float w1[64]; float b0, b1, ls0, ls1; };

// decoder block, second half: cross-attention (query = obs_rep) + proj + LN2(rep + .) + MLP + LN3;
// block 0 then emits QKV1 of block 1, block 1 the logits head and the per-token log-prob / entropy.
template <int RPT>
__global__ void __launch_bounds__(NT, 2)
dec_attn2(PolicyBufs b, TrainBufs tb, int blk, const float* __restrict__ actions, const float* __restrict__ ava,
          float* __restrict__ logp, float* __restrict__ entropy, long n_tok, int L, int n_disc, int depth) {
    constexpr int TM = 16 * RPT;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem s = carve<RPT>(smem_raw, depth);
    const int tid = threadIdx.x, rg = tid >> 4, cg = tid & 15;
    const long tile0 = (long)blockIdx.x * TM;
    const int n_valid = (int)min((long)TM, n_tok - tile0);
    Ring ring;
    ring.init(s.ring, s.bars, b.w, c_seq_dec_attn2[blk], blk == 0 ? 6 : 4, blk == 0 ? 6 : 4, depth);
    token_rows<RPT>(s, tile0, n_tok, L, b.K2[blk], b.V2[blk], true);
    load_tile<RPT>(s.sA, b.Q2[blk] + tile0 * 64, 64, n_valid);
    if (blk == 1) for (int i = tid; i < lay::DEC_H3_SIZE; i += NT) s.sSmall[i] = b.w[lay::DEC_H3 + i];
    __syncthreads();
    ring.prefetch();
    attention_rows<RPT, true>(s.sA, s.sB, s.kptr, s.vptr, s.nt, tb.on ? tb.d_lse2[blk] + tile0 * 2 : nullptr, n_valid);
    SaveOut sv1 = {nullptr, nullptr, nullptr}, sv2 = sv1, sv3 = sv1, sv4 = sv1;
    if (tb.on) {
        store_tile<RPT>(tb.d_y2[blk] + tile0 * 64, 64, s.sB, n_valid);
        sv1.xhat = tb.d_xh2[blk] + tile0 * 64; sv1.rstd = tb.d_rs2[blk] + tile0;
        sv2.pre = tb.d_m0[blk] + tile0 * 64;
        sv3.xhat = tb.d_xh3[blk] + tile0 * 64; sv3.rstd = tb.d_rs3[blk] + tile0;
        sv4.pre = tb.d_hh + tile0 * 64; sv4.xhat = tb.d_xhh + tile0 * 64; sv4.rstd = tb.d_rsh + tile0;
    }
    stage<EP_RES | EP_LN, RPT>(ring, s.sB, b.rep + tile0 * 64, 64, n_valid, s.sX, XS, TM, rg, cg, &sv1, n_valid);
    __syncthreads(); ring.prefetch();
    stage<EP_GELU, RPT>(ring, s.sX, s.sA, XS, TM, rg, cg, &sv2, n_valid);
    __syncthreads(); ring.prefetch();
    stage<EP_RES | EP_LN, RPT>(ring, s.sA, s.sX, XS, TM, s.sX, XS, TM, rg, cg, &sv3, n_valid);
    __syncthreads(); ring.prefetch();
    if (blk == 0) {
        store_tile<RPT>(b.X + tile0 * 64, 64, s.sX, n_valid);
        stage<0, RPT>(ring, s.sX, (tb.on ? tb.d_q1[1] : b.Q) + tile0 * 64, 64, n_valid, rg, cg);
        stage<0, RPT>(ring, s.sX, b.K1[1] + tile0 * 64, 64, n_valid, rg, cg);
        stage<0, RPT>(ring, s.sX, b.V1[1] + tile0 * 64, 64, n_valid, rg, cg);
    } else {
        stage<EP_GELU | EP_LN, RPT>(ring, s.sX, s.sA, XS, TM, rg, cg, &sv4, n_valid);
        __syncthreads();
        if (tid < n_valid) {
            const long g = tile0 + tid;
            const int pos = (int)(g % L);
            const float l0 = row_dot(s.sA, tid, s.sSmall) + s.sSmall[lay::DEC_H3_B];
            const float l1 = row_dot(s.sA, tid, s.sSmall + 64) + s.sSmall[lay::DEC_H3_B + 1];
            const float a = actions[g];
            if (pos < n_disc) {
                float z0, z1;
                masked_logz(l0, l1, ava ? ava + g * 2 : nullptr, z0, z1);
                logp[g] = a != 0.f ? z1 : z0;
                const float c0 = fmaxf(z0, -3.4028234663852886e38f), c1 = fmaxf(z1, -3.4028234663852886e38f);
                entropy[g] = -(expf(c0) * c0 + expf(c1) * c1);
            } else {
                const float sd = 0.5f / (1.0f + expf(-s.sSmall[lay::DEC_H3_LOGSTD + 1]));
                const float d = a - l1;
                logp[g] = -(d * d) / (2.0f * sd * sd) - logf(sd) - HALF_LOG_2PI;
                entropy[g] = 0.5f + HALF_LOG_2PI + logf(sd);
            }
        }
    }
}

// -------------------------------------------------------------------------------------------------
// persistent KV-cached decode: one CTA owns 64 envs at a time and walks the schedule.
// -------------------------------------------------------------------------------------------------
// phase profiler: thread 0 of CTA 0 adds the cycles since the previous marker to slot K
#define PHASE(K)                                                                     \
    do {                                                                             \
        if (a.phase_cycles && blockIdx.x == 0 && threadIdx.x == 0) {                 \
            const long long now_ = clock64();                                        \
            atomicAdd(a.phase_cycles + (K), (unsigned long long)(now_ - t_mark));    \
            t_mark = now_;                                                           \
        }                                                                            \
    } while (0)

template <int RPT>
__global__ void __launch_bounds__(NT, 2)
decode_kernel(PolicyBufs b, DecodeArgs a, int depth) {
    constexpr int TM = 16 * RPT;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem s = carve<RPT>(smem_raw, depth);
    const int tid = threadIdx.x, rg = tid >> 4, cg = tid & 15;
    const int L = a.L;
    const long n_tiles = (a.B + TM - 1) / TM;
    long my_tiles = 0;
    for (long t = blockIdx.x; t < n_tiles; t += gridDim.x) ++my_tiles;
    Ring ring;
    ring.init(s.ring, s.bars, b.w, c_seq_decode, lay::DEC_FOLD_STEP_BLOBS, (int)(my_tiles * a.n_steps * lay::DEC_FOLD_STEP_BLOBS), depth);
    float* sEmb = s.sSmall;            // [6][64]
    float* sHead = s.sSmall + 384;     // DEC_H3 block
    int* sCode = reinterpret_cast<int*>(s.sSmall + 640);   // [TM]
    build_emb_table(sEmb, b.w);
    for (int i = tid; i < lay::DEC_H3_SIZE; i += NT) sHead[i] = b.w[lay::DEC_H3 + i];
    __syncthreads();
    ring.prefetch();

    long long t_mark = clock64();
    const long slot0 = (long)blockIdx.x * TM;
    const long slot_stride = 4L * L * 64;      // [blk 2][attn 2][L][64] cached attention inputs per env slot
    float* kv0 = b.kv + slot0 * slot_stride;
    const int nd = a.mode == MATAS_MODE_SEMI ? L - 1 : L;
    const float sd1 = 0.5f / (1.0f + expf(-sHead[lay::DEC_H3_LOGSTD + 1]));

    for (long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long env0 = tile * TM;
        const int n_valid = (int)min((long)TM, a.B - env0);
        for (int i = tid; i < TM * 8; i += NT) s.bits[i] = 0u;
        DecSched sched;
        sched.init(L, a.mode, a.deterministic, a.stride);
        DecStep st;
        __syncthreads();
        while (sched.next(st)) {
            const int j = st.j;
            // ---- inputs of this position: token embedding, obs_rep row -----------------------------
            if (tid < TM) {
                int code;
                const uint32_t* bt = s.bits + tid * 8;
                if (st.kind == 1) code = j == 0 ? 1 : 0;
                else if (st.kind == 0) code = j == 0 ? 1 : 2 + (int)((bt[(j - 1) >> 5] >> ((j - 1) & 31)) & 1u);
                else code = (j == 0 ? 4 : 2) + (int)((bt[j >> 5] >> (j & 31)) & 1u);
                sCode[tid] = code;
            }
            __syncthreads();
            for (int idx = tid; idx < TM * 16; idx += NT) {
                const int r = idx >> 4, c = idx & 15;
                *reinterpret_cast<float4*>(s.sX + r * XS + c * 4) = *reinterpret_cast<const float4*>(sEmb + sCode[r] * 64 + c * 4);
            }
            __syncthreads();
            PHASE(0);
            for (int blk = 0; blk < 2; ++blk) {
                float* cx1 = kv0 + (blk * 2 + 0) * (long)L * 64;     // self-attention inputs (block input x)
                float* cx2 = kv0 + (blk * 2 + 1) * (long)L * 64;     // cross-attention key/value inputs (x after LN1)
                // ---- self-attention: cache x_j, folded queries, attend over the cached inputs -----------
                store_tile<RPT>(cx1 + (long)j * 64, slot_stride, s.sX, n_valid);
                if (a.prefetch && blk == 0) prefetch_cache_rows<RPT>(cx1, slot_stride, n_valid, j);
                stage<0, RPT>(ring, s.sX, s.sA, XS, TM, rg, cg);
                stage<0, RPT>(ring, s.sX, s.sB, XS, TM, rg, cg);
                if (tid < TM) {
                    s.kptr[tid] = cx1 + tid * slot_stride;
                    s.nt[tid] = tid < n_valid ? j + 1 : 0;
                }
                __syncthreads(); ring.prefetch();
                PHASE(1);
                attention_absorbed<RPT>(s.sA, s.sB, s.kptr, s.nt);
                PHASE(2);
                if (a.prefetch) prefetch_cache_rows<RPT>(cx2, slot_stride, n_valid, j);
                stage2<EP_RES | EP_LN, RPT>(ring, s.sA, s.sB, s.sX, XS, TM, s.sX, XS, TM, rg, cg);
                __syncthreads(); ring.prefetch();
                PHASE(3);
                // ---- cross-attention: cache x1_j, queries q2' precomputed from obs_rep --------------------
                store_tile<RPT>(cx2 + (long)j * 64, slot_stride, s.sX, n_valid);
                load_tile<RPT>(s.sA, b.Q2p[blk][0] + (env0 * L + j) * 64, (long)L * 64, n_valid);
                load_tile<RPT>(s.sB, b.Q2p[blk][1] + (env0 * L + j) * 64, (long)L * 64, n_valid);
                if (tid < TM) s.kptr[tid] = cx2 + tid * slot_stride;
                __syncthreads();
                PHASE(4);
                attention_absorbed<RPT>(s.sA, s.sB, s.kptr, s.nt);
                PHASE(5);
                stage2<EP_RES | EP_LN, RPT>(ring, s.sA, s.sB, b.rep + (env0 * L + j) * 64, (long)L * 64, n_valid, s.sX, XS, TM, rg, cg);
                __syncthreads(); ring.prefetch();
                PHASE(6);
                if (a.prefetch && blk == 0) prefetch_cache_rows<RPT>(kv0 + 2 * (long)L * 64, slot_stride, n_valid, j);
                stage<EP_GELU, RPT>(ring, s.sX, s.sA, XS, TM, rg, cg);
                __syncthreads(); ring.prefetch();
                PHASE(7);
                stage<EP_RES | EP_LN, RPT>(ring, s.sA, s.sX, XS, TM, s.sX, XS, TM, rg, cg);
                __syncthreads(); ring.prefetch();
                PHASE(8);
            }
            // ---- logits head + action pick ---------------------------------------------------------
            if (st.emit) {
                stage<EP_GELU | EP_LN, RPT>(ring, s.sX, s.sA, XS, TM, rg, cg);
            } else {
                ring.acquire(); ring.release();
            }
            __syncthreads(); ring.prefetch();
            PHASE(9);
            if (st.emit && tid < n_valid) {
                const long env = env0 + tid;
                const long g = env * L + j;
                const float l0 = row_dot(s.sA, tid, sHead) + sHead[lay::DEC_H3_B];
                const float l1 = row_dot(s.sA, tid, sHead + 64) + sHead[lay::DEC_H3_B + 1];
                if (j < nd) {
                    float z0, z1;
                    masked_logz(l0, l1, a.ava ? a.ava + g * 2 : nullptr, z0, z1);
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
            __syncthreads();
            PHASE(10);
        }
    }
}

template <typename K>
int set_smem(K kernel, size_t bytes) {
    CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return MATAS_OK;
}

}  // namespace

// -------------------------------------------------------------------------------------------------
// host launchers
// -------------------------------------------------------------------------------------------------
// Tile geometry: 32-row tiles (2 rows per thread) with two CTAs per SM whenever both fit in shared
// memory -- one CTA's attention (memory latency) then overlaps the other's GEMMs -- else 64-row tiles.
static int sm_count() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) n = 148;
    }
    return n;
}

int policy_launch_geometry(int L, long rows, int* rpt, int* depth, size_t* bytes, int kind) {
    if (L > 256) {
        matas_set_error("n_agent %d > 256 is not supported (per-env action bitset)", L);
        return MATAS_ERR_LIMIT;
    }
    // Two CTAs per SM in every tile kernel (one CTA's attention / barriers overlap the other's GEMMs).
    // kind 0: token-parallel passes; kind 1: decode.  MATAS_RPT_TOKEN / MATAS_RPT_DECODE (2 | 4) override the
    // rows-per-thread choice for tuning: 4 -> 64-row tiles (4x4 outputs per thread), 2 -> 32-row tiles.
    // 64-row tiles give the better GEMM (4x4 outputs per thread) but need enough tiles to fill the resident
    // CTA slots; smaller problems use 32-row tiles (measured: C3 decode 9.8 vs 11.0 ms, 4096 x 101 13.6 vs 20.2 ms).
    const char* ov = getenv(kind == 0 ? "MATAS_RPT_TOKEN" : "MATAS_RPT_DECODE");
    const long tiles64 = (rows + 63) / 64;
    const long want = kind == 0 ? 2L * sm_count() : (3L * sm_count()) / 2;
    *rpt = ov ? atoi(ov) : (tiles64 >= want ? 4 : 2);
    if (*rpt != 2 && *rpt != 4) *rpt = 4;
    *depth = 3;
    *bytes = smem_bytes(16 * *rpt, *depth);
    return MATAS_OK;
}

int policy_decode_geometry(int L, long B, int n_sm, int* grid, long* slots) {
    if (policy_use_tc()) {
        if (L > 256) { matas_set_error("n_agent %d > 256 is not supported (per-env action bitset)", L); return MATAS_ERR_LIMIT; }
        return decode_tc_geometry(B, n_sm, grid, slots);
    }
    int rpt, depth; size_t bytes;
    int rc = policy_launch_geometry(L, B, &rpt, &depth, &bytes, 1);
    if (rc) return rc;
    const int tm = 16 * rpt;
    const long n_tiles = (B + tm - 1) / tm;
    const long resident = (long)n_sm * 2;
    *grid = (int)(n_tiles < resident ? n_tiles : resident);
    *slots = (long)*grid * tm;
    return MATAS_OK;
}

int policy_decode_steps(int L, int mode, int deterministic, int stride) {
    DecSched s; s.init(L, mode, deterministic, stride);
    DecStep st; int n = 0;
    while (s.next(st)) ++n;
    return n;
}

int policy_init_constants() { return MATAS_OK; }

static TrainBufs train_or_off(const TrainBufs* tb) {
    TrainBufs t;
    if (tb) t = *tb; else { memset(&t, 0, sizeof(t)); }
    return t;
}

#define GEOMETRY(L, ROWS, KIND)                                       \
    int rpt, depth; size_t bytes;                                     \
    { int rc_ = policy_launch_geometry(L, ROWS, &rpt, &depth, &bytes, KIND); if (rc_) return rc_; } \
    const int tm = 16 * rpt
// launch KERNEL<..., RPT> for the chosen geometry
#define LAUNCH(KERNEL2, KERNEL4, GRID, ...)                           \
    do {                                                              \
        int rc_;                                                      \
        if (rpt == 2) { if ((rc_ = set_smem(KERNEL2, bytes))) return rc_; KERNEL2<<<GRID, NT, bytes, st>>>(__VA_ARGS__); } \
        else { if ((rc_ = set_smem(KERNEL4, bytes))) return rc_; KERNEL4<<<GRID, NT, bytes, st>>>(__VA_ARGS__); }          \
    } while (0)

// MATAS_TC=0 keeps the dense stages of the token-parallel kernels on the fp32 SIMT path (A/B measurements)
bool policy_use_tc() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("MATAS_TC"); v = e ? atoi(e) : 1; }
    return v != 0;
}

int launch_encoder(const PolicyBufs& b, const float* obs, float* values, long B, int L, int obs_dim, cudaStream_t st,
                   const TrainBufs* tbp, int for_decode) {
    const TrainBufs tb = train_or_off(tbp);
    GEOMETRY(L, B * L, 0);
    const long n_tok = B * L;
    const unsigned grid = (unsigned)((n_tok + tm - 1) / tm);
    if (policy_use_tc()) {
        int rc_ = launch_enc_embed_tc(b, tb, obs, n_tok, obs_dim, st);
        return rc_ ? rc_ : launch_enc_blocks_tc(b, tb, values, n_tok, L, for_decode, st);
    }
    LAUNCH(enc_embed_qkv<2>, enc_embed_qkv<4>, grid, b, tb, obs, n_tok, obs_dim, depth);
    LAUNCH((enc_block<0, 2>), (enc_block<0, 4>), grid, b, tb, values, n_tok, L, depth, 0);
    LAUNCH((enc_block<1, 2>), (enc_block<1, 4>), grid, b, tb, values, n_tok, L, depth, for_decode);
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}

int launch_q2_from_rep(const PolicyBufs& b, const float* rep_in, long B, int L, cudaStream_t st, int for_decode) {
    GEOMETRY(L, B * L, 0);
    const long n_tok = B * L;
    const unsigned grid = (unsigned)((n_tok + tm - 1) / tm);
    LAUNCH(q2_from_rep<2>, q2_from_rep<4>, grid, b, rep_in, n_tok, depth, for_decode);
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}

int launch_teacher_forced(const PolicyBufs& b, const float* actions, const float* ava, float* logp, float* entropy,
                          long B, int L, int mode, cudaStream_t st, const TrainBufs* tbp) {
    const TrainBufs tb = train_or_off(tbp);
    GEOMETRY(L, B * L, 0);
    const long n_tok = B * L;
    const unsigned grid = (unsigned)((n_tok + tm - 1) / tm);
    const int n_disc = mode == MATAS_MODE_SEMI ? L - 1 : L;
    if (policy_use_tc()) return launch_teacher_forced_tc(b, tb, actions, ava, logp, entropy, n_tok, L, n_disc, st);
    LAUNCH(dec_embed_qkv<2>, dec_embed_qkv<4>, grid, b, tb, actions, n_tok, L, depth);
    for (int blk = 0; blk < 2; ++blk) {
        LAUNCH(dec_attn1<2>, dec_attn1<4>, grid, b, tb, blk, n_tok, L, depth);
        LAUNCH(dec_attn2<2>, dec_attn2<4>, grid, b, tb, blk, actions, ava, logp, entropy, n_tok, L, n_disc, depth);
    }
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}

int launch_decode(const PolicyBufs& b, const DecodeArgs& a, int grid, cudaStream_t st) {
    if (policy_use_tc()) {
        if (a.phase_cycles) {
            matas_set_error("the decode phase profiler instruments the SIMT kernel: set MATAS_TC=0");
            return MATAS_ERR_UNSUPPORTED;
        }
        return launch_decode_tc(b, a, grid, sm_count(), st);
    }
    GEOMETRY(a.L, a.B, 1);
    (void)tm;
    LAUNCH(decode_kernel<2>, decode_kernel<4>, grid, b, a, depth);
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}
