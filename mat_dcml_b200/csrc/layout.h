// Packed device weight layout of the MAT-AS policy (n_embd 64, n_head 2, n_block 2, obs_dim <= 8,
// act_dim 2).  The flat parameter vector the C-ABI takes is in the reference's state_dict order
// (mat_src/mat/algorithms/mat/algorithm/ma_transformer.py:119-203); matas_policy_set_weights
// repacks it into "blobs", one per 64x64 Linear, each holding everything its fused GEMM stage reads:
//   [ W^T  (in-major, 64x64) | bias 64 | LayerNorm weight 64 | LayerNorm bias 64 ]  = 4288 floats
// so that one cp.async.bulk brings a whole stage into shared memory.
#pragma once

namespace lay {
constexpr int D = 64;
constexpr int BLOB = 4288;                 // floats; 17152 bytes, a multiple of 16
constexpr int BLOB_BYTES = BLOB * 4;
constexpr int B_BIAS = 4096, B_LNW = 4160, B_LNB = 4224;

// encoder blobs, usage order inside a block
enum { EQ = 0, EK, EV, EP, EM0, EM2, ENC_PER_BLOCK };
// decoder per-step blobs, usage order inside a block (Q2 is applied to obs_rep ahead of the decode)
enum { DQ1 = 0, DK1, DV1, DP1, DK2, DV2, DP2, DM0, DM2, DEC_PER_BLOCK };
constexpr int DEC_STEP_BLOBS = 2 * DEC_PER_BLOCK + 1;   // + head.0

// small parameter groups
constexpr int ENC_EMB = 0;                 // ln_w[8] ln_b[8] Wobs^T[8][64] b[64] enc.ln w[64] b[64]
constexpr int ENC_EMB_LNW = 0, ENC_EMB_LNB = 8, ENC_EMB_WT = 16, ENC_EMB_B = 528, ENC_EMB_LN2W = 592,
              ENC_EMB_LN2B = 656, ENC_EMB_SIZE = 720;
constexpr int DEC_EMB = ENC_EMB + ENC_EMB_SIZE;   // Wa^T[4][64] (row 3 zero) dec.ln w[64] b[64]
constexpr int DEC_EMB_WT = 0, DEC_EMB_LNW = 256, DEC_EMB_LNB = 320, DEC_EMB_SIZE = 384;
constexpr int ENC_H3 = DEC_EMB + DEC_EMB_SIZE;    // w[64] b[1]
constexpr int ENC_H3_SIZE = 80;
constexpr int DEC_H3 = ENC_H3 + ENC_H3_SIZE;      // w0[64] w1[64] b[2] log_std[2]
constexpr int DEC_H3_B = 128, DEC_H3_LOGSTD = 130, DEC_H3_SIZE = 144;
constexpr int SMALL_SIZE = DEC_H3 + DEC_H3_SIZE;  // 1328 floats

constexpr int ENC_BLOBS = SMALL_SIZE;                               // [2][ENC_PER_BLOCK] then head.0
constexpr int ENC_H0 = ENC_BLOBS + 2 * ENC_PER_BLOCK * BLOB;
constexpr int DEC_STEP = ENC_H0 + BLOB;                             // DEC_STEP_BLOBS blobs in usage order
constexpr int DEC_Q2 = DEC_STEP + DEC_STEP_BLOBS * BLOB;            // [2] blobs
constexpr int TOTAL = DEC_Q2 + 2 * BLOB;
// backward region: the same 34 Linear weights, untransposed ([out][in] = the [k][n] layout of dx = dy.W)
constexpr int N_BLOBS = (TOTAL - ENC_BLOBS) / BLOB;       // 34
constexpr int BW_BASE = TOTAL;
constexpr int BW_MAT = 4096;
constexpr int BW_BYTES = BW_MAT * 4;
__host__ __device__ constexpr int bw(int blob_off) { return BW_BASE + ((blob_off - ENC_BLOBS) / BLOB) * BW_MAT; }
// decode-only folded blobs ("absorbed" attention: the cache holds the 64-float layer INPUT of each
// attention instead of its K and V rows).  Per decoder block, per head h:
//   FA_h : q'_h = scale * Wk_h^T (Wq_h x + bq_h)        self-attention query map          (x -> 64)
//   FB_h : out += (Wp[:, h] Wv_h) ctx_h   (+ Wp bv + bp and LN1 in FB_1)   self-attention output map
//   FC_h : the same for cross-attention (+ LN2 in FC_1)
//   FQ_h : q2'_h = scale * Wk2_h^T (Wq2_h rep + bq2_h)   cross-attention query map, applied ahead of the decode
enum { FA0 = 0, FA1, FB0, FB1, FC0, FC1, FQ0, FQ1, FOLD_PER_BLOCK };
constexpr int FOLD_BASE = BW_BASE + N_BLOBS * BW_MAT;
__host__ __device__ constexpr int fold_blob(int blk, int which) { return FOLD_BASE + (blk * FOLD_PER_BLOCK + which) * BLOB; }
constexpr int DEC_FOLD_STEP_BLOBS = 2 * 8 + 1;        // per block: FA0 FA1 FB0 FB1 FC0 FC1 M0 M2 ; + head.0
constexpr int TOTAL_ALL = FOLD_BASE + 2 * FOLD_PER_BLOCK * BLOB;
// tensor-core ("half") blobs: for every forward / folded blob at float offset `off` the same 17152 bytes at
// hb(off) hold the weight as tcgen05 B operands — W[n][k] in fp16, K-major, SWIZZLE_128B (row n at n*128 B,
// 16-byte chunk index XOR (n % 8)): the hi part fp16(w) (8 KB), then the lo part fp16((w - hi) * 2^11) (8 KB) —
// followed by the same fp32 bias | LN weight | LN bias tail (so B_BIAS / B_LNW / B_LNB index it unchanged).
constexpr int HB_BASE = TOTAL_ALL;
__host__ __device__ constexpr int hb(int blob_off) { return HB_BASE + (blob_off - ENC_BLOBS); }
constexpr int TOTAL_HB = HB_BASE + (TOTAL_ALL - ENC_BLOBS);
static_assert((HB_BASE - ENC_BLOBS) % 4 == 0 && ENC_BLOBS % 4 == 0, "half blobs must stay 16-byte aligned for cp.async.bulk");
constexpr int HB_LO_BYTES = 8192;                     // byte offset of the lo part inside a half blob

__host__ __device__ constexpr int enc_blob(int blk, int which) { return ENC_BLOBS + (blk * ENC_PER_BLOCK + which) * BLOB; }
__host__ __device__ constexpr int dec_blob(int blk, int which) { return DEC_STEP + (blk * DEC_PER_BLOCK + which) * BLOB; }
constexpr int DEC_H0 = DEC_STEP + 2 * DEC_PER_BLOCK * BLOB;
__host__ __device__ constexpr int dec_q2(int blk) { return DEC_Q2 + blk * BLOB; }
}  // namespace lay
