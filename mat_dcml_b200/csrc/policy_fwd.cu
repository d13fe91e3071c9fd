// Host launchers of the MAT-AS policy forward passes for sm_100a, and the one token-parallel kernel that is
// not in policy_tc.cu: the cross-attention queries from a caller-supplied obs_rep.
//
//   encoder      : enc_embed_qkv_tc -> {attn_tc -> enc_block_tc<k>} x 2            (policy_tc.cu, attn_tc.cu)
//   evaluate     : encoder + dec_embed_qkv_tc -> {attn_tc, dec_attn1_tc, attn_tc, dec_attn2_tc} x 2
//   get_actions  : encoder + decode_tc_kernel                                        (decode_tc.cu)
//
// Reference semantics: mat_src/mat/algorithms/mat/algorithm/ma_transformer.py:24-230 and
// mat_src/mat/algorithms/utils/transformer_act.py:8-189.
#include <stdlib.h>
#include <string.h>

#include "policy_fwd.h"
#include "tile.cuh"

using namespace tile;

namespace {

constexpr int Q2_DEPTH = 3;
constexpr size_t Q2_SMEM = ((size_t)Q2_DEPTH * lay::BLOB + TILE_F) * 4 + 8 * 8 + 16;

// plain / folded cross-attention query maps of the two decoder blocks (float offsets into the packed weights)
__constant__ int c_seq_q2[2] = {lay::dec_q2(0), lay::dec_q2(1)};
__constant__ int c_seq_q2p[4] = {lay::fold_blob(0, lay::FQ0), lay::fold_blob(0, lay::FQ1), lay::fold_blob(1, lay::FQ0),
                                 lay::fold_blob(1, lay::FQ1)};

// cross-attention queries from a caller-supplied obs_rep (the transformer_act entry points start
// from obs_rep, transformer_act.py:30,103,131,176): rep_in -> b.rep copy, Q2[blk] = attn2.query(rep)
// (teacher forcing) or the folded q2'_h (decode).  64-token tiles, fp32 SIMT stages (tile.cuh): two to four
// [64 x 64] products per tile whose time is the token rows' HBM traffic.
__global__ void __launch_bounds__(NT, 2)
q2_from_rep(PolicyBufs b, const float* __restrict__ rep_in, long n_tok, int for_decode) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* s_ring = reinterpret_cast<float*>(smem_raw);
    float* sX = s_ring + (size_t)Q2_DEPTH * lay::BLOB;
    uint64_t* bars = reinterpret_cast<uint64_t*>((reinterpret_cast<uintptr_t>(sX + TILE_F) + 15) & ~(uintptr_t)15);
    const int tid = threadIdx.x, rg = tid >> 4, cg = tid & 15;
    const long tile0 = (long)blockIdx.x * TM;
    const int n_valid = (int)min((long)TM, n_tok - tile0);
    Ring ring;
    if (for_decode) ring.init(s_ring, bars, b.w, c_seq_q2p, 4, 4, Q2_DEPTH);
    else ring.init(s_ring, bars, b.w, c_seq_q2, 2, 2, Q2_DEPTH);
    load_tile(sX, rep_in + tile0 * 64, 64, n_valid);
    __syncthreads();
    ring.prefetch();
    if (rep_in != b.rep) {
        for (int idx = tid; idx < TM * 16; idx += NT) {
            const int r = idx >> 4, c = idx & 15;
            if (r < n_valid) *reinterpret_cast<float4*>(b.rep + (tile0 + r) * 64 + c * 4) = *reinterpret_cast<const float4*>(sX + r * XS + c * 4);
        }
    }
    if (for_decode) {
        stage<0>(ring, sX, b.Q2p[0][0] + tile0 * 64, 64, n_valid, rg, cg);
        stage<0>(ring, sX, b.Q2p[0][1] + tile0 * 64, 64, n_valid, rg, cg);
        __syncthreads(); ring.prefetch();
        stage<0>(ring, sX, b.Q2p[1][0] + tile0 * 64, 64, n_valid, rg, cg);
        stage<0>(ring, sX, b.Q2p[1][1] + tile0 * 64, 64, n_valid, rg, cg);
    } else {
        stage<0>(ring, sX, b.Q2[0] + tile0 * 64, 64, n_valid, rg, cg);
        stage<0>(ring, sX, b.Q2[1] + tile0 * 64, 64, n_valid, rg, cg);
    }
}

}  // namespace

// -------------------------------------------------------------------------------------------------
// host launchers
// -------------------------------------------------------------------------------------------------
int policy_check_agents(int L) {
    if (L > 256) {
        matas_set_error("n_agent %d > 256 is not supported (per-env action bitset, attention key tiles)", L);
        return MATAS_ERR_LIMIT;
    }
    return MATAS_OK;
}

int policy_decode_geometry(int L, long B, int n_sm, int* grid, long* slots, int* rows) {
    int rc = policy_check_agents(L);
    if (rc) return rc;
    return decode_tc_geometry(B, n_sm, grid, slots, rows);
}

int policy_decode_steps(int L, int mode, int deterministic, int stride) {
    DecSched s; s.init(L, mode, deterministic, stride);
    DecStep st; int n = 0;
    while (s.next(st)) ++n;
    return n;
}

static TrainBufs train_or_off(const TrainBufs* tb) {
    TrainBufs t;
    if (tb) t = *tb; else { memset(&t, 0, sizeof(t)); }
    return t;
}

int launch_encoder(const PolicyBufs& b, const float* obs, float* values, long B, int L, int obs_dim, int n_sm, cudaStream_t st,
                   const TrainBufs* tbp, int for_decode) {
    int rc = policy_check_agents(L);
    if (rc) return rc;
    const TrainBufs tb = train_or_off(tbp);
    const long n_tok = B * L;
    if ((rc = launch_enc_embed_tc(b, tb, obs, n_tok, obs_dim, st))) return rc;
    return launch_enc_blocks_tc(b, tb, values, n_tok, L, for_decode, n_sm, st);
}

int launch_q2_from_rep(const PolicyBufs& b, const float* rep_in, long B, int L, cudaStream_t st, int for_decode) {
    int rc = policy_check_agents(L);
    if (rc) return rc;
    const long n_tok = B * L;
    CUDA_TRY(cudaFuncSetAttribute(q2_from_rep, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Q2_SMEM));
    q2_from_rep<<<(unsigned)((n_tok + TM - 1) / TM), NT, Q2_SMEM, st>>>(b, rep_in, n_tok, for_decode);
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}

int launch_teacher_forced(const PolicyBufs& b, const float* actions, const float* ava, float* logp, float* entropy,
                          long B, int L, int mode, int n_sm, cudaStream_t st, const TrainBufs* tbp) {
    int rc = policy_check_agents(L);
    if (rc) return rc;
    const TrainBufs tb = train_or_off(tbp);
    const int n_disc = mode == MATAS_MODE_SEMI ? L - 1 : L;
    return launch_teacher_forced_tc(b, tb, actions, ava, logp, entropy, B * L, L, n_disc, n_sm, st);
}
