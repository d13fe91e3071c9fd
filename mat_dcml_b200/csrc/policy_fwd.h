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
    float* Q2[2];        // attn2.query(obs_rep) per decoder block
    float* rep;          // encoder output obs_rep
    float* kv;           // decode K/V cache: [grid*64 env slots][blk 2][attn 2][k|v][L][64]
};

struct DecodeArgs {
    long B;
    int L, mode, deterministic, stride, n_steps;
    uint64_t seed;
    uint32_t step;
    long env_offset;
    const float* ava;    // [B][L][2] or null
    const float* u;      // [B][L] injected uniforms or null (Philox)
    const float* z;      // [B][2] injected normals or null (Philox)
    float* actions;      // [B][L]
    float* logp;         // [B][L]
};

struct matas_policy;
struct PolicyView { int n_agent, mode, device, n_sm, weights_set; const float* packed; };
int matas_policy_view(const matas_policy* p, PolicyView* v);

int policy_launch_geometry(int L, int* depth, int* lpad, size_t* bytes);
int policy_decode_steps(int L, int mode, int deterministic, int stride);
int policy_init_constants();
int launch_encoder(const PolicyBufs& b, const float* obs, float* values, long B, int L, int obs_dim, cudaStream_t st);
int launch_q2_from_rep(const PolicyBufs& b, const float* rep_in, long B, int L, cudaStream_t st);
int launch_teacher_forced(const PolicyBufs& b, const float* actions, const float* ava, float* logp, float* entropy,
                          long B, int L, int mode, cudaStream_t st);
int launch_decode(const PolicyBufs& b, const DecodeArgs& a, int grid, cudaStream_t st);
