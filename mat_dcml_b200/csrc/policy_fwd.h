// Internal interface between the C-ABI layer (cabi.cu) and the policy kernels (policy_fwd.cu).
#pragma once
#include "common.cuh"

#define MATAS_MODE_SEMI 0       /* L-1 two-way categorical agents + 1 Gaussian agent (DCML) */
#define MATAS_MODE_DISCRETE 1   /* all agents two-way categorical */

// device buffers of one policy handle; token buffers are [n_tok][64] fp32, n_tok = B*L
struct PolicyBufs {
    const float* w;      // packed weights (layout.h)
    float* X;            // residual stream
    float* Q;            // queries of the block being processed
    float* K[2];         // encoder keys / values per block
    float* V[2];
    float* K1[2];        // teacher-forced decoder: self-attention keys / values per block
    float* V1[2];
    float* K2[2];        // teacher-forced decoder: cross-attention keys / values per block
    float* V2[2];
    float* Q2[2];        // attn2.query(obs_rep) per decoder block (teacher-forced passes)
    float* Q2p[2][2];    // decode: folded cross-attention queries q2'_h per block and head
    float* rep;          // encoder output obs_rep
    float* Y;            // attention output of the block being processed (register-tiled attention kernel, attn_tile.cu)
    float* kv;           // decode cache: [env slots][blk 2][attn 2][L][64] layer inputs of each attention
    const float* x1tab;  // block-0 cross-attention inputs by (position, ones so far, previous action) or null (decode_tc.cu)
    const float* x1spec; // the same for the speculative rows of the deterministic schedule's stride blocks, or null
};

// training-mode saves of the teacher-forced forward + work buffers of the backward; [n_tok][64] fp32
// unless noted.  xh = LayerNorm-normalised value before the affine map, rs = 1/std per token.
struct TrainBufs {
    int on;                                              // 0: inference (nothing is saved)
    float *e_h0, *e_xh0, *e_rs0, *e_x0;                  // encoder embed: pre-GELU, LN, output
    float *d_x0;                                         // decoder embed output
    float *e_q[2], *e_y[2], *e_lse[2];                   // per block: queries, attention output, lse [tok][2]
    float *e_xh1[2], *e_rs1[2], *e_m0[2], *e_xh2[2], *e_rs2[2];
    float *e_hh, *e_xhh, *e_rsh;                         // value head: pre-GELU, LN
    float *d_q1[2], *d_y1[2], *d_lse1[2], *d_xh1[2], *d_rs1[2];
    float *d_y2[2], *d_lse2[2], *d_xh2[2], *d_rs2[2], *d_m0[2], *d_xh3[2], *d_rs3[2];
    float *d_hh, *d_xhh, *d_rsh;                         // logits head
    float *gS[9];                                        // backward: three rotating sets of three layer-output gradient buffers handed to
                                                         // the weight-gradient kernel (a set is reused only after its jobs were launched)
    float *gX, *gRep, *gYa, *gYb, *gDa, *gDb;            // backward: residual-stream / obs_rep / attention-output grads, D [tok][2]
};

struct DecodeArgs {
    long B;
    int L, mode, deterministic, stride, n_steps;
    int rows;            // envs per CTA tile (decode_tc_geometry); the cache scratch holds grid * rows env slots
    uint64_t seed;
    uint32_t step;
    long env_offset;
    const float* ava;    // [B][L][2] or null
    const float* u;      // [B][L] injected uniforms or null (Philox)
    const float* z;      // [B][2] injected normals or null (Philox)
    float* actions;      // [B][L]
    float* logp;         // [B][L]
};

constexpr float MASK_LOGIT = -1e10f;
constexpr float HALF_LOG_2PI = 0.91893853320467274178f;

// log-softmax of the two (availability-masked) logits of a discrete agent (transformer_act.py:13-22)
__device__ __forceinline__ void masked_logz(float l0, float l1, const float* ava2, float& z0, float& z1) {
    if (ava2) {
        if (ava2[0] == 0.f) l0 = MASK_LOGIT;
        if (ava2[1] == 0.f) l1 = MASK_LOGIT;
    }
    const float m = fmaxf(l0, l1);
    const float lse = m + logf(expf(l0 - m) + expf(l1 - m));
    z0 = l0 - lse;
    z1 = l1 - lse;
}

// -------------------------------------------------------------------------------------------------
// autoregressive decode schedule (transformer_act.py:37-98 and :138-172), shared by host and device
// -------------------------------------------------------------------------------------------------
struct DecStep { int j; int kind; int emit; };   // kind 0: true predecessor, 1: speculative zero row, 2: own action

struct DecSched {
    int L, nd, stride, mode, det;
    int s, e, j, phase, go, started;
    __host__ __device__ void init(int L_, int mode_, int det_, int stride_) {
        L = L_; mode = mode_; det = det_; stride = stride_ < 1 ? 1 : stride_;
        nd = mode == MATAS_MODE_SEMI ? L - 1 : L;
        s = 0; e = 1; j = 0; phase = 0; go = 1; started = 0;
    }
    __host__ __device__ bool next(DecStep& st) {
        if (!det) {
            if (j >= L) return false;
            st.j = j++; st.kind = 0; st.emit = 1;
            return true;
        }
        for (;;) {
            if (phase == 0) {           // entering a block
                if (!(e <= L && go)) return false;
                if (e == L) go = 0;
                phase = 1; j = s;
            }
            if (phase == 1) {           // speculative pass over [s, e)
                if (j < e) {
                    st.j = j; st.emit = 1;
                    st.kind = mode == MATAS_MODE_SEMI ? (j == s ? 0 : 1) : 1;
                    ++j;
                    return true;
                }
                const bool more = mode == MATAS_MODE_SEMI ? (go != 0) : (e < L);
                if (!more) return false;
                phase = 2; j = mode == MATAS_MODE_SEMI ? s + 1 : s;
            }
            if (phase == 2) {           // commit pass: recompute with the decided actions
                if (j < e) {
                    st.j = j; st.emit = 0; st.kind = mode == MATAS_MODE_SEMI ? 0 : 2;
                    ++j;
                    return true;
                }
                if (mode == MATAS_MODE_SEMI) {
                    if (e < nd) { s = e; e = min(e + stride, nd); }
                    else { s = e; e = min(e + 1, L); }
                } else {
                    s = e; e = min(e + 4, L);
                }
                phase = 0;
            }
        }
    }
};

struct matas_policy;
struct PolicyView { int n_agent, mode, device, n_sm, weights_set; const float* packed; const float* x1tab; };
int matas_policy_view(const matas_policy* p, PolicyView* v);

int policy_check_agents(int L);
// decode launch geometry for B envs on n_sm SMs: CTAs, env slots of cache scratch (grid * rows), envs per CTA tile
int policy_decode_geometry(int L, long B, int n_sm, int* grid, long* slots, int* rows);
int policy_decode_steps(int L, int mode, int deterministic, int stride);
int launch_encoder(const PolicyBufs& b, const float* obs, float* values, long B, int L, int obs_dim, int n_sm, cudaStream_t st,
                   const TrainBufs* tb = nullptr, int for_decode = 0);
int launch_q2_from_rep(const PolicyBufs& b, const float* rep_in, long B, int L, cudaStream_t st, int for_decode);
int launch_teacher_forced(const PolicyBufs& b, const float* actions, const float* ava, float* logp, float* entropy,
                          long B, int L, int mode, int n_sm, cudaStream_t st, const TrainBufs* tb = nullptr);
// tensor-core (tcgen05) kernels: policy_tc.cu, attn_tc.cu, decode_tc.cu
int decode_tc_geometry(long B, int n_sm, int* grid, long* slots, int* rows);
int launch_decode_tc(const PolicyBufs& b, const DecodeArgs& a, int grid, cudaStream_t st);
int launch_x1_table(const float* packed, float* x1tab, int L, cudaStream_t st);
long decode_tc_spec_rows(int L, int stride);
int launch_x1_spec_table(const float* packed, float* x1spec, int L, int stride, cudaStream_t st);
bool attn_tile_supported(int L);      // n_agent <= 64: the register-tiled attention backward (attn_bwd_tile.cu)
int launch_attn_tc(const float* Q, const float* K, const float* V, float* Y, float* lse, long n_env, int L, int causal,
                   int n_sm, cudaStream_t st);
int launch_teacher_forced_tc(const PolicyBufs& b, const TrainBufs& tb, const float* actions, const float* ava, float* logp,
                             float* entropy, long n_tok, int L, int n_disc, int n_sm, cudaStream_t st);
int launch_enc_embed_tc(const PolicyBufs& b, const TrainBufs& tb, const float* obs, long n_tok, int obs_dim, cudaStream_t st);
int launch_enc_blocks_tc(const PolicyBufs& b, const TrainBufs& tb, float* values, long n_tok, int L, int for_decode, int n_sm,
                         cudaStream_t st);
