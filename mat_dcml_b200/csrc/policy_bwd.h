// Internal interface of the evaluate backward (policy_bwd.cu).
#pragma once
#include "layout.h"
#include "policy_fwd.h"

// offsets of every trainable tensor inside the flat parameter / gradient vector (state_dict order)
struct GradMap {
    int w[lay::N_BLOBS], b[lay::N_BLOBS];          // the 34 64x64 Linear layers, indexed like the forward blobs
    int lnw[lay::N_BLOBS], lnb[lay::N_BLOBS];      // LayerNorm fused after that Linear (-1: none)
    int enc_ln7_w, enc_ln7_b, enc_wobs, enc_bobs, enc_ln_w, enc_ln_b, enc_h3_w, enc_h3_b;
    int log_std, dec_wa, dec_ln_w, dec_ln_b, dec_h3_w, dec_h3_b;
};

int launch_evaluate_backward(const PolicyBufs& b, const TrainBufs& t, const GradMap& gm, float* grad, const float* obs,
                             const float* actions, const float* ava, const float* d_values, const float* d_logp,
                             const float* d_ent, long M, int L, int obs_dim, int mode, int n_sm, cudaStream_t st);

// ---- weight gradients on the tensor cores (bwd_dw_tc.cu) ------------------------------------------------
enum { DW_X_PLAIN = 0, DW_X_AFFINE = 1, DW_X_GELU = 2 };     // how the layer input X is rebuilt from what the forward saved
struct DwJob {
    const float* dY;     // [tok][64] gradient w.r.t. the layer output (scratch written by the fused backward kernels)
    const float* X;      // [tok][64] layer input as saved: the value itself, x_hat (AFFINE) or the pre-activation (GELU)
    int recipe;
    int ln;              // AFFINE: packed offset of the forward blob holding the LayerNorm weight | bias
    int gW;              // offset of the [out][in] weight gradient in the flat gradient
};
struct DwJobs { DwJob j[8]; int n; };
int launch_bwd_dw(const float* packed, float* grad, const DwJobs& jobs, long n_tok, int n_sm, cudaStream_t st);

// register-tiled per-env attention backward (attn_bwd_tile.cu), n_agent <= 64
int launch_attn_bwd_tile(const float* Q, const float* K, const float* V, const float* gY, const float* lse, const float* D, float* dQ,
                         float* dK, float* dV, long n_env, int L, int causal, cudaStream_t st);

// input gradient of the q / k / v projections + their bias gradients on the tensor cores (bwd_dx_tc.cu)
struct DxArgs {
    const float *dq, *dk, *dv;   // [tok][64] from attn_bwd_tile.cu
    int blob[3];                 // forward blob offsets (layout.h) of the q / k / v Linear layers
    int gb[3];                   // bias-gradient offsets in the flat gradient
    float* gX;                   // kind 0: += dq.Wq + dk.Wk + dv.Wv ; kind 1: = dk.Wk + dv.Wv
    float* gRep;                 // kind 1: += dq.Wq
};
int launch_bwd_qkv_dx(const float* packed, float* grad, const DxArgs& a, long n_tok, int kind, int n_sm, cudaStream_t st);

// [LN_F bwd -> MLP bwd ->] LN_M bwd -> residual grad -> projection bwd (encoder block tail, decoder block halves)
struct TailArgs {
    const float* gOut;     // grad wrt the block output
    const float *xhF, *rsF, *m0, *xhM, *rsM, *Y;
    int lnF, lnM;          // packed offsets of the forward blobs that hold these LayerNorm params
    int seq[3];            // backward weight offsets: W2, W0, Wp   (with_mlp) or Wp only
    int gW2, gb2, gW0, gb0, gWp, gbp, glnFw, glnFb, glnMw, glnMb;   // flat-gradient offsets
    float* gRes; int res_add;
    float *gY, *gD;
    int with_mlp;
    int fW0;               // forward blob offset of mlp.0 (mlp.2 and proj share lnF / lnM)
    float* sY[3];          // non-null: write dS / dM0 / d(proj out) there instead of accumulating dW2 / dW0 / dWp here
};

// the same chain with its three dY.W products on the tensor cores (bwd_mlp_tc.cu); requires a.sY
int launch_bwd_mlp_tc(const float* packed, float* grad, const TailArgs& a, long n_tok, int n_sm, cudaStream_t st);
