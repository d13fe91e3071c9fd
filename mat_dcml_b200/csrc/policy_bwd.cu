// Backward pass of the teacher-forced MAT-AS forward (MultiAgentTransformer.forward,
// mat_src/mat/algorithms/mat/algorithm/ma_transformer.py:257-295) for sm_100a.  The reference gets
// these gradients from torch autograd inside MATTrainer.ppo_update (mat_trainer.py:146-147); here
// they are hand-written kernels that accumulate into ONE flat fp32 gradient vector laid out like
// the flat parameter vector (state_dict order), which is also what the multi-GPU all-reduce sees.
//
// Structure: persistent CTAs walk 64-token tiles; weight gradients (dW = dY^T X) live in registers
// across tiles and are flushed with atomics once per CTA; dX = dY W streams the untransposed
// weights through the same TMA ring as the forward.  Attention backward recomputes probabilities
// from the saved Q/K and log-sum-exp (query side for dQ, key side for dK/dV; D = rowsum(dY*Y)).
#include <stdlib.h>
#include <string.h>

#include "policy_bwd.h"
#include "tile.cuh"

using namespace tile;

namespace {

constexpr float MASK_LOGIT = -1e10f;
constexpr float SCALE = 0.17677669529663687f;   // 1/sqrt(32)
constexpr int BDEPTH = 3;          // ring depth of the 1-CTA/SM kernels
constexpr int ADEPTH = 2;          // ring depth of bwd_attn (2 CTAs/SM)
constexpr int NTILES = 4;
constexpr float LOG2E = 1.4426950408889634f;

struct BSmem {
    float* ring;      // depth * BLOB
    float* T[NTILES]; // tiles [TM][XS]
    float* sSmall;    // 1024 floats
    uint64_t* bars;
};

__host__ __device__ inline size_t bsmem_bytes(int depth) {
    const size_t f = (size_t)depth * lay::BLOB + NTILES * TILE_F + 1024;
    return f * 4 + 8 * 8 + 32;
}

__device__ inline BSmem bcarve(unsigned char* base, int depth) {
    BSmem s;
    float* f = reinterpret_cast<float*>(base);
    s.ring = f; f += (size_t)depth * lay::BLOB;
    for (int i = 0; i < NTILES; ++i) { s.T[i] = f; f += TILE_F; }
    s.sSmall = f; f += 1024;
    uintptr_t p = (reinterpret_cast<uintptr_t>(f) + 15) & ~(uintptr_t)15;
    s.bars = reinterpret_cast<uint64_t*>(p);
    return s;
}

// ---- register fragments: thread (rg, cg) owns rows 4rg..4rg+3, cols 4cg..4cg+3 of a tile --------
struct Frag { float v[4][4]; };

__device__ __forceinline__ void frag_load_g(Frag& f, const float* __restrict__ g, long row0, int n_valid, int rg, int cg) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int r = rg * 4 + i;
        float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < n_valid) x = __ldcg(reinterpret_cast<const float4*>(g + (row0 + r) * 64 + cg * 4));
        f.v[i][0] = x.x; f.v[i][1] = x.y; f.v[i][2] = x.z; f.v[i][3] = x.w;
    }
}
__device__ __forceinline__ void frag_store_g(const Frag& f, float* __restrict__ g, long row0, int n_valid, int rg, int cg, bool add) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int r = rg * 4 + i;
        if (r < n_valid) {
            float4* p = reinterpret_cast<float4*>(g + (row0 + r) * 64 + cg * 4);
            float4 x = make_float4(f.v[i][0], f.v[i][1], f.v[i][2], f.v[i][3]);
            if (add) { const float4 o = *p; x.x += o.x; x.y += o.y; x.z += o.z; x.w += o.w; }
            *p = x;
        }
    }
}
__device__ __forceinline__ void frag_store_s(const Frag& f, float* __restrict__ t, int rg, int cg) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
        *reinterpret_cast<float4*>(t + (rg * 4 + i) * XS + cg * 4) = make_float4(f.v[i][0], f.v[i][1], f.v[i][2], f.v[i][3]);
}
__device__ __forceinline__ void frag_load_s(Frag& f, const float* __restrict__ t, int rg, int cg) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float4 x = *reinterpret_cast<const float4*>(t + (rg * 4 + i) * XS + cg * 4);
        f.v[i][0] = x.x; f.v[i][1] = x.y; f.v[i][2] = x.z; f.v[i][3] = x.w;
    }
}
__device__ __forceinline__ float4 ld4(const float* p, int cg) { return *reinterpret_cast<const float4*>(p + cg * 4); }
__device__ __forceinline__ float f4(const float4& a, int j) { return j == 0 ? a.x : (j == 1 ? a.y : (j == 2 ? a.z : a.w)); }

// column-sum accumulator (bias / LayerNorm parameter gradients): 4 columns per thread
struct Col { float v[4]; __device__ void zero() { v[0] = v[1] = v[2] = v[3] = 0.f; } };
__device__ __forceinline__ void col_add(Col& c, const Frag& f) {
#pragma unroll
    for (int j = 0; j < 4; ++j) c.v[j] += f.v[0][j] + f.v[1][j] + f.v[2][j] + f.v[3][j];
}
__device__ __forceinline__ void col_add_mul(Col& c, const Frag& a, const Frag& b) {
#pragma unroll
    for (int j = 0; j < 4; ++j) c.v[j] += a.v[0][j] * b.v[0][j] + a.v[1][j] * b.v[1][j] + a.v[2][j] * b.v[2][j] + a.v[3][j] * b.v[3][j];
}
__device__ __forceinline__ void col_flush(const Col& c, float* __restrict__ grad, int off, int cg) {
    if (off < 0) return;
#pragma unroll
    for (int j = 0; j < 4; ++j) atomicAdd(grad + off + cg * 4 + j, c.v[j]);
}

// dX fragment = dY(tile in shared) . W (untransposed weights from the ring)
__device__ __forceinline__ void gemm_dx(Ring& ring, const float* __restrict__ dY, Frag& out, int rg, int cg, bool accumulate) {
    const float* W = ring.acquire();
    float acc[4][4];
    gemm(dY, W, acc, rg, cg);
    ring.release();
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) out.v[i][j] = accumulate ? out.v[i][j] + acc[i][j] : acc[i][j];
}

// y = xhat * w + b (recompute a LayerNorm output from the saved normalised value)
__device__ __forceinline__ void ln_affine(Frag& y, const Frag& xh, const float* __restrict__ lnw, const float* __restrict__ lnb, int cg) {
    const float4 w = ld4(lnw, cg), b = ld4(lnb, cg);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) y.v[i][j] = xh.v[i][j] * f4(w, j) + f4(b, j);
}

// LayerNorm backward on a fragment.  dy -> dx (in place); accumulates d(weight), d(bias).
__device__ __forceinline__ void ln_bwd(Frag& d, const Frag& xh, const float* __restrict__ rstd_g, long row0, int n_valid,
                                       const float* __restrict__ lnw, Col& gw, Col& gb, int rg, int cg) {
    col_add_mul(gw, d, xh);
    col_add(gb, d);
    const float4 w = ld4(lnw, cg);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int r = rg * 4 + i;
        float g[4], s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) { g[j] = d.v[i][j] * f4(w, j); s1 += g[j]; s2 += g[j] * xh.v[i][j]; }
        s1 = half_warp_sum(s1) * (1.0f / 64.0f);
        s2 = half_warp_sum(s2) * (1.0f / 64.0f);
        const float rs = r < n_valid ? rstd_g[row0 + r] : 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) d.v[i][j] = rs * (g[j] - s1 - xh.v[i][j] * s2);
    }
}

__device__ __forceinline__ float gelu_grad(float h) {
    const float cdf = 0.5f * (1.0f + erf_as(h * 0.70710678118654752f));       // A&S 7.1.26: 1.5e-7 absolute
    return cdf + h * 0.3989422804014327f * expf(-0.5f * h * h);
}
__device__ __forceinline__ void gelu_bwd(Frag& d, const Frag& h) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) d.v[i][j] *= gelu_grad(h.v[i][j]);
}
// -------------------------------------------------------------------------------------------------
// attention backward.  KIND 0: q, k, v all project the same input (encoder attention, decoder
// self-attention); KIND 1: decoder cross-attention (q from obs_rep, k/v from the decoder stream).
// -------------------------------------------------------------------------------------------------
struct AttnArgs {
    const float *Q, *K, *V, *lse, *gY, *gD;     // [tok][64] / [tok][2]
    const float *xh, *rs;                       // saved LN whose output is the k/v (and, KIND 0, q) input
    int ln;                                     // packed offset of the blob holding that LN (-1: input is `xin` as is)
    const float* xin;                           // KIND 0 with ln < 0, or KIND 1 query input (obs_rep)
    int seq[3];                                 // backward weights Wq, Wk, Wv
    int gWq, gbq, gWk, gbk, gWv, gbv;
    float* gX;                                  // KIND 0: += ; KIND 1: = (grad wrt the k/v input)
    float* gRep;                                // KIND 1: += dq.Wq
    int causal, L;
    float* sY[3];                               // non-null: write dq / dk / dv there instead of accumulating dWq / dWk / dWv here
    int pre;                                    // dq / dk / dv were already computed into sY by attn_bwd_tile.cu: only the projections run here
};

// Sum a and b over the 8 lanes of a head (lane bit 2 = `up`) with 4 shuffles instead of 6: the two
// halves of the group reduce one value each, then swap the totals.
__device__ __forceinline__ void head_sum2(float& a, float& b, unsigned mask, bool up) {
    float keep = up ? b : a;
    const float send = up ? a : b;
    keep += __shfl_xor_sync(mask, send, 4);
    keep += __shfl_xor_sync(mask, keep, 2);
    keep += __shfl_xor_sync(mask, keep, 1);
    const float other = __shfl_xor_sync(mask, keep, 4);
    a = up ? other : keep;
    b = up ? keep : other;
}
__device__ __forceinline__ float dot4(const float4& a, const float4& b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
__device__ __forceinline__ void axpy4(float4& acc, float p, const float4& v) {
    acc.x = fmaf(p, v.x, acc.x); acc.y = fmaf(p, v.y, acc.y); acc.z = fmaf(p, v.z, acc.z); acc.w = fmaf(p, v.w, acc.w);
}

// Attention backward, one half-warp per row of the tile (lane l: columns 4l..4l+3, lanes 0-7 head 0,
// lanes 8-15 head 1), every global row read coalesced; probabilities are recomputed from the saved
// log-sum-exp, so there is no score buffer and no CTA barrier inside.
//   query side, row = query i:  dq_i = scale * sum_t p_it (dY_i.v_t - D_i) k_t             t in [0, hi)
//   key side,   row = key j  :  dv_j = sum_i p_ij dY_i ;  dk_j = scale * sum_i p_ij (dY_i.v_j - D_i) q_i   i in [lo, L)
// with p = exp(scale q.k - lse) per head.
template <int KB>
__device__ __forceinline__ void attn_bwd_rows(const AttnArgs& a, long row0, long n_tok, float* __restrict__ sDq,
                                              float* __restrict__ sDv, float* __restrict__ sDk) {
    const int tid = threadIdx.x, hw = tid >> 4, l = tid & 15, h = l >> 3, L = a.L;
    const bool up = (l & 4) != 0;
    const unsigned hmask = 0xffffu << (tid & 16);
    const float c2 = SCALE * LOG2E;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = hw; r < TM; r += NT / 16) {
        const long g = row0 + r;
        float4 dq = zero4, dv = zero4, dk = zero4;
        if (g < n_tok) {
            const long env = g / L, base = env * L;
            const int pos = (int)(g - base);
            const float4* Qe = reinterpret_cast<const float4*>(a.Q + base * 64) + l;     // row t of the env at [t * 16]
            const float4* Ke = reinterpret_cast<const float4*>(a.K + base * 64) + l;
            const float4* Ve = reinterpret_cast<const float4*>(a.V + base * 64) + l;
            const float4* Ye = reinterpret_cast<const float4*>(a.gY + base * 64) + l;
            const float* lse_e = a.lse + base * 2 + h;
            const float* D_e = a.gD + base * 2 + h;
            {   // ---- query side
                const int hi = a.causal ? pos + 1 : L;
                const float4 q = __ldg(Qe + pos * 16), dy = __ldg(Ye + pos * 16);
                const float lse2 = __ldg(lse_e + pos * 2) * LOG2E, Dh = __ldg(D_e + pos * 2);
                for (int t0 = 0; t0 < hi; t0 += KB) {
                    float4 k[KB], v[KB];
#pragma unroll
                    for (int i = 0; i < KB; ++i) {
                        const int t = min(t0 + i, hi - 1);
                        k[i] = __ldg(Ke + t * 16); v[i] = __ldg(Ve + t * 16);
                    }
#pragma unroll
                    for (int i = 0; i < KB; ++i) {
                        float s1 = dot4(q, k[i]), s2 = dot4(dy, v[i]);
                        head_sum2(s1, s2, hmask, up);
                        const float p = t0 + i < hi ? exp2f(fmaf(s1, c2, -lse2)) : 0.f;
                        axpy4(dq, p * (s2 - Dh), k[i]);
                    }
                }
            }
            {   // ---- key side
                const int lo = a.causal ? pos : 0;
                const float4 k = __ldg(Ke + pos * 16), v = __ldg(Ve + pos * 16);
                for (int i0 = lo; i0 < L; i0 += KB) {
                    float4 q[KB], dy[KB];
                    float ls[KB], dd[KB];
#pragma unroll
                    for (int i = 0; i < KB; ++i) {
                        const int t = min(i0 + i, L - 1);
                        q[i] = __ldg(Qe + t * 16); dy[i] = __ldg(Ye + t * 16);
                        ls[i] = __ldg(lse_e + t * 2); dd[i] = __ldg(D_e + t * 2);
                    }
#pragma unroll
                    for (int i = 0; i < KB; ++i) {
                        float s1 = dot4(q[i], k), s2 = dot4(dy[i], v);
                        head_sum2(s1, s2, hmask, up);
                        const float p = i0 + i < L ? exp2f(fmaf(s1, c2, -ls[i] * LOG2E)) : 0.f;
                        axpy4(dv, p, dy[i]);
                        axpy4(dk, p * (s2 - dd[i]), q[i]);
                    }
                }
            }
        }
        *reinterpret_cast<float4*>(sDq + r * XS + l * 4) = make_float4(dq.x * SCALE, dq.y * SCALE, dq.z * SCALE, dq.w * SCALE);
        *reinterpret_cast<float4*>(sDv + r * XS + l * 4) = dv;
        *reinterpret_cast<float4*>(sDk + r * XS + l * 4) = make_float4(dk.x * SCALE, dk.y * SCALE, dk.z * SCALE, dk.w * SCALE);
    }
}

template <int KIND>
__global__ void __launch_bounds__(NT, 2)
bwd_attn(const float* __restrict__ packed, float* __restrict__ grad, AttnArgs a, long n_tok) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    BSmem s = bcarve(smem_raw, ADEPTH);
    const int tid = threadIdx.x, rg = tid >> 4, cg = tid & 15;
    const long n_tiles = (n_tok + TM - 1) / TM;
    long mine = 0;
    for (long t = blockIdx.x; t < n_tiles; t += gridDim.x) ++mine;
    int* sseq = reinterpret_cast<int*>(s.sSmall + 512);
    if (tid < 3) sseq[tid] = a.seq[tid];
    if (a.ln >= 0) for (int i = tid; i < 128; i += NT) s.sSmall[i] = packed[a.ln + lay::B_LNW + i];
    Ring ring;
    ring.init(s.ring, s.bars, packed, sseq, 3, (int)(mine * 3), ADEPTH, lay::BW_BYTES);
    __syncthreads();
    ring.prefetch();
    float *sDq = s.T[1], *sDv = s.T[2], *sDk = s.T[3];
    Col dbq, dbk, dbv; dbq.zero(); dbk.zero(); dbv.zero();
    for (long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long row0 = tile * TM;
        const int n_valid = (int)min((long)TM, n_tok - row0);
        if (a.pre) {
            load_tile(sDq, a.sY[0] + row0 * 64, 64, n_valid);
            load_tile(sDk, a.sY[1] + row0 * 64, 64, n_valid);
            load_tile(sDv, a.sY[2] + row0 * 64, 64, n_valid);
        } else {
            attn_bwd_rows<4>(a, row0, n_tok, sDq, sDv, sDk);
        }
        // ---- projections: dW, db, dX -------------------------------------------------------------
        Frag dx;
        const bool put = !a.pre;                                         // hand dq / dk / dv to the weight-gradient kernel (bwd_dw_tc.cu)
        if (KIND == 1) {
            __syncthreads();
            { Frag f; frag_load_s(f, sDq, rg, cg); col_add(dbq, f); if (put) frag_store_g(f, a.sY[0], row0, n_valid, rg, cg, false); }
            gemm_dx(ring, sDq, dx, rg, cg, false);
            frag_store_g(dx, a.gRep, row0, n_valid, rg, cg, true);
            __syncthreads();
            ring.prefetch();
        }
        __syncthreads();
        if (KIND == 0) {
            { Frag f; frag_load_s(f, sDq, rg, cg); col_add(dbq, f); if (put) frag_store_g(f, a.sY[0], row0, n_valid, rg, cg, false); }
            gemm_dx(ring, sDq, dx, rg, cg, false);
            __syncthreads();
            ring.prefetch();
        }
        { Frag f; frag_load_s(f, sDk, rg, cg); col_add(dbk, f); if (put) frag_store_g(f, a.sY[1], row0, n_valid, rg, cg, false); }
        gemm_dx(ring, sDk, dx, rg, cg, KIND == 0);
        __syncthreads();
        ring.prefetch();
        { Frag f; frag_load_s(f, sDv, rg, cg); col_add(dbv, f); if (put) frag_store_g(f, a.sY[2], row0, n_valid, rg, cg, false); }
        gemm_dx(ring, sDv, dx, rg, cg, true);
        frag_store_g(dx, a.gX, row0, n_valid, rg, cg, KIND == 0);       // KIND 0: += ; KIND 1: =
        __syncthreads();
        ring.prefetch();
    }
    col_flush(dbq, grad, a.gbq, cg); col_flush(dbk, grad, a.gbk, cg); col_flush(dbv, grad, a.gbv, cg);
}

// -------------------------------------------------------------------------------------------------
// heads.  DEC: (d_logp, d_entropy) -> logits head backward -> gX (grad of the last decoder block output)
//         ENC: d_values -> value head backward -> gRep += (grad of obs_rep)
// -------------------------------------------------------------------------------------------------
struct HeadArgs {
    const float *d_a, *d_b;          // DEC: d_logp, d_entropy [tok];  ENC: d_values [tok], unused
    const float *actions, *ava;      // DEC only
    const float *hh, *xhh, *rsh;     // saved pre-GELU, LN xhat, rstd
    const float *xh_in, *rs_in;      // DEC: LN3 of block 1 (its output is the head input); ENC: null (input = rep)
    const float* rep;                // ENC head input
    int ln_in;                       // DEC: packed blob offset holding LN3 params
    int blob_h0;                     // forward blob of head.0 (holds head.2 LN params); backward weight via lay::bw
    int h3;                          // packed offset of the head.3 group
    int gW0, gb0, glnw, glnb, gW3, gb3, glogstd;
    float* gOut;                     // DEC: gX (store); ENC: gRep (add)
    int L, n_disc;
    float* sY;                       // non-null: write d(head.0 output) there instead of accumulating dW0 here
};

template <int DEC>
__global__ void __launch_bounds__(NT, 1)
bwd_head(const float* __restrict__ packed, float* __restrict__ grad, HeadArgs a, long n_tok) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    BSmem s = bcarve(smem_raw, BDEPTH);
    const int tid = threadIdx.x, rg = tid >> 4, cg = tid & 15;
    const long n_tiles = (n_tok + TM - 1) / TM;
    long mine = 0;
    for (long t = blockIdx.x; t < n_tiles; t += gridDim.x) ++mine;
    int* sseq = reinterpret_cast<int*>(s.sSmall + 512);
    if (tid == 0) sseq[0] = lay::bw(a.blob_h0);
    for (int i = tid; i < 128; i += NT) {
        s.sSmall[i] = packed[a.blob_h0 + lay::B_LNW + i];                  // head.2 LN w | b
        if (DEC) s.sSmall[128 + i] = packed[a.ln_in + lay::B_LNW + i];     // LN3 w | b
    }
    for (int i = tid; i < (DEC ? lay::DEC_H3_SIZE : lay::ENC_H3_SIZE); i += NT) s.sSmall[256 + i] = packed[a.h3 + i];
    Ring ring;
    ring.init(s.ring, s.bars, packed, sseq, 1, (int)mine, BDEPTH, lay::BW_BYTES);
    __syncthreads();
    ring.prefetch();
    const float* w3 = s.sSmall + 256;     // DEC: w0[64] w1[64] b[2] log_std[2];  ENC: w[64] b[1]
    float* sdl = s.sSmall + 640;          // [TM][2] d(logits) / d(value)
    Col db0, glw, glb, gw3a, gw3b; db0.zero(); glw.zero(); glb.zero(); gw3a.zero(); gw3b.zero();
    float gb3a = 0.f, gb3b = 0.f, gls = 0.f;
    for (long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long row0 = tile * TM;
        const int n_valid = (int)min((long)TM, n_tok - row0);
        Frag xh, nh;
        frag_load_g(xh, a.xhh, row0, n_valid, rg, cg);
        ln_affine(nh, xh, s.sSmall, s.sSmall + 64, cg);
        frag_store_s(nh, s.T[0], rg, cg);
        __syncthreads();
        if (tid < TM) {
            float dl0 = 0.f, dl1 = 0.f;
            if (tid < n_valid) {
                const long g = row0 + tid;
                if (DEC) {
                    float l0 = row_dot(s.T[0], tid, w3) + w3[lay::DEC_H3_B];
                    float l1 = row_dot(s.T[0], tid, w3 + 64) + w3[lay::DEC_H3_B + 1];
                    const float dlp = a.d_a[g], dent = a.d_b[g], act = a.actions[g];
                    const int pos = (int)(g % a.L);
                    if (pos < a.n_disc) {
                        bool m0 = false, m1 = false;
                        if (a.ava) { m0 = a.ava[g * 2] == 0.f; m1 = a.ava[g * 2 + 1] == 0.f; }
                        if (m0) l0 = MASK_LOGIT;
                        if (m1) l1 = MASK_LOGIT;
                        const float mx = fmaxf(l0, l1);
                        const float lse = mx + logf(expf(l0 - mx) + expf(l1 - mx));
                        const float z0 = l0 - lse, z1 = l1 - lse, p0 = expf(z0), p1 = expf(z1);
                        const float c0 = fmaxf(z0, -3.4028234663852886e38f), c1 = fmaxf(z1, -3.4028234663852886e38f);
                        const float H = -(p0 * c0 + p1 * c1);
                        const int ai = act != 0.f;
                        // d logp/dl_c = [c==a] - p_c ;  dH/dl_c = -p_c (z_c + H)
                        dl0 = dlp * ((ai == 0 ? 1.f : 0.f) - p0) + dent * (-p0 * (c0 + H));
                        dl1 = dlp * ((ai == 1 ? 1.f : 0.f) - p1) + dent * (-p1 * (c1 + H));
                        if (m0) dl0 = 0.f;       // masked logits were overwritten: no gradient reaches them
                        if (m1) dl1 = 0.f;
                    } else {
                        const float sg = 1.0f / (1.0f + expf(-w3[lay::DEC_H3_LOGSTD + 1]));
                        const float sd = 0.5f * sg;
                        const float dd = act - l1;
                        dl1 = dlp * dd / (sd * sd);
                        gls += (dlp * (dd * dd / (sd * sd) - 1.0f) + dent) * (1.0f - sg);
                    }
                } else {
                    dl0 = a.d_a[g];
                }
            }
            sdl[tid * 2] = dl0; sdl[tid * 2 + 1] = dl1;
            gb3a += dl0; gb3b += dl1;
        }
        __syncthreads();
        // d(head.3 weight) and d(nh)
        Frag d;
        const float4 wa = ld4(w3, cg), wb = DEC ? ld4(w3 + 64, cg) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = rg * 4 + i;
            const float a0 = sdl[r * 2], a1 = sdl[r * 2 + 1];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                gw3a.v[j] += a0 * nh.v[i][j];
                gw3b.v[j] += a1 * nh.v[i][j];
                d.v[i][j] = a0 * f4(wa, j) + a1 * f4(wb, j);
            }
        }
        ln_bwd(d, xh, a.rsh, row0, n_valid, s.sSmall, glw, glb, rg, cg);
        Frag h;
        frag_load_g(h, a.hh, row0, n_valid, rg, cg);
        gelu_bwd(d, h);                                                    // d = d(pre-GELU of head.0)
        col_add(db0, d);
        frag_store_s(d, s.T[1], rg, cg);
        frag_store_g(d, a.sY, row0, n_valid, rg, cg, false);               // weight gradient: bwd_dw_tc.cu
        __syncthreads();
        Frag dx;
        gemm_dx(ring, s.T[1], dx, rg, cg, false);
        frag_store_g(dx, a.gOut, row0, n_valid, rg, cg, DEC ? false : true);
        __syncthreads();
        ring.prefetch();
    }
    col_flush(db0, grad, a.gb0, cg); col_flush(glw, grad, a.glnw, cg); col_flush(glb, grad, a.glnb, cg);
    col_flush(gw3a, grad, a.gW3, cg);
    if (DEC) col_flush(gw3b, grad, a.gW3 + 64, cg);
    if (tid < TM) {
        atomicAdd(grad + a.gb3, gb3a);
        if (DEC) { atomicAdd(grad + a.gb3 + 1, gb3b); if (a.glogstd >= 0) atomicAdd(grad + a.glogstd + 1, gls); }
    }
}

// -------------------------------------------------------------------------------------------------
// embeddings
// -------------------------------------------------------------------------------------------------
struct EmbArgs {
    const float* gX;                  // grad wrt the embedding output
    const float *h0, *xh0, *rs0;      // ENC saved; DEC recomputed from the action code
    const float* obs; int obs_dim;    // ENC
    const float* actions; int L;      // DEC
    int gln_w, gln_b, gW, gb, gln7_w, gln7_b;
};

// encoder: x0 = LN(GELU(W_obs . LN7(obs) + b))      (ma_transformer.py:133-136,148-151)
__global__ void __launch_bounds__(NT, 1)
bwd_embed_enc(const float* __restrict__ packed, float* __restrict__ grad, EmbArgs a, long n_tok) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    BSmem s = bcarve(smem_raw, BDEPTH);
    const int tid = threadIdx.x, rg = tid >> 4, cg = tid & 15;
    const long n_tiles = (n_tok + TM - 1) / TM;
    for (int i = tid; i < lay::ENC_EMB_SIZE; i += NT) s.sSmall[i] = packed[lay::ENC_EMB + i];
    __syncthreads();
    float* sO = s.T[1];       // [TM][8] LN7 output
    float* sOh = s.T[2];      // [TM][8] LN7 normalised
    const int od = a.obs_dim;
    Col glw, glb, gb; glw.zero(); glb.zero(); gb.zero();
    float accW[2] = {0.f, 0.f};       // dW_obs entries (c, k) = (idx/8, idx%8), idx = tid, tid+256
    float acc7w = 0.f, acc7b = 0.f;   // tid < 8: LN7 weight / bias grads
    for (long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long row0 = tile * TM;
        const int n_valid = (int)min((long)TM, n_tok - row0);
        if (tid < TM) {
            float v[8], mean = 0.f;
            for (int k = 0; k < 8; ++k) { v[k] = (k < od && tid < n_valid) ? a.obs[(row0 + tid) * od + k] : 0.f; mean += v[k]; }
            mean /= (float)od;
            float var = 0.f;
            for (int k = 0; k < od; ++k) var += (v[k] - mean) * (v[k] - mean);
            const float rs = rsqrtf(var / (float)od + 1e-5f);
            for (int k = 0; k < 8; ++k) {
                const float xh = k < od ? (v[k] - mean) * rs : 0.f;
                sOh[tid * 8 + k] = xh;
                sO[tid * 8 + k] = k < od ? xh * s.sSmall[lay::ENC_EMB_LNW + k] + s.sSmall[lay::ENC_EMB_LNB + k] : 0.f;
            }
        }
        Frag d, xh, h;
        frag_load_g(d, a.gX, row0, n_valid, rg, cg);
        frag_load_g(xh, a.xh0, row0, n_valid, rg, cg);
        ln_bwd(d, xh, a.rs0, row0, n_valid, s.sSmall + lay::ENC_EMB_LN2W, glw, glb, rg, cg);
        frag_load_g(h, a.h0, row0, n_valid, rg, cg);
        gelu_bwd(d, h);
        col_add(gb, d);
        frag_store_s(d, s.T[0], rg, cg);
        __syncthreads();
        for (int q = 0; q < 2; ++q) {
            const int idx = tid + q * NT, c = idx >> 3, k = idx & 7;
            float acc = 0.f;
            for (int r = 0; r < n_valid; ++r) acc += s.T[0][r * XS + c] * sO[r * 8 + k];
            accW[q] += acc;
        }
        // LN7 parameter grads: do7[r][k] = sum_c dh0[r][c] Wobs[c][k]
        for (int idx = tid; idx < TM * 8; idx += NT) {
            const int r = idx >> 3, k = idx & 7;
            float acc = 0.f;
            if (r < n_valid && k < od) for (int c = 0; c < 64; ++c) acc += s.T[0][r * XS + c] * s.sSmall[lay::ENC_EMB_WT + k * 64 + c];
            s.T[3][idx] = acc;
        }
        __syncthreads();
        if (tid < 8) for (int r = 0; r < n_valid; ++r) { acc7w += s.T[3][r * 8 + tid] * sOh[r * 8 + tid]; acc7b += s.T[3][r * 8 + tid]; }
        __syncthreads();
    }
    col_flush(glw, grad, a.gln_w, cg); col_flush(glb, grad, a.gln_b, cg); col_flush(gb, grad, a.gb, cg);
    for (int q = 0; q < 2; ++q) {
        const int idx = tid + q * NT, c = idx >> 3, k = idx & 7;
        if (k < od) atomicAdd(grad + a.gW + c * od + k, accW[q]);
    }
    if (tid < od) { atomicAdd(grad + a.gln7_w + tid, acc7w); atomicAdd(grad + a.gln7_b + tid, acc7b); }
}

// decoder: x0 = LN(GELU(W_a . shifted))   (ma_transformer.py:224-225); shifted row = start flag | one-hot(prev action)
__global__ void __launch_bounds__(NT, 1)
bwd_embed_dec(const float* __restrict__ packed, float* __restrict__ grad, EmbArgs a, long n_tok) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    BSmem s = bcarve(smem_raw, BDEPTH);
    const int tid = threadIdx.x, rg = tid >> 4, cg = tid & 15;
    const long n_tiles = (n_tok + TM - 1) / TM;
    // tables for the 3 reachable token codes (1 start, 2 action 0, 3 action 1): pre-GELU a, xhat, rstd
    float* tA = s.sSmall;             // [4][64]
    float* tXh = s.sSmall + 256;      // [4][64]
    float* tRs = s.sSmall + 512;      // [4]
    float* lnw = s.sSmall + 520;      // [64]
    int* sCode = reinterpret_cast<int*>(s.sSmall + 600);
    if (tid < 64) {
        const int c = tid >> 4, cg2 = tid & 15;                 // code index c -> code c+1 ... use c in 0..3 (0 unused)
        const float* Wt = packed + lay::DEC_EMB + lay::DEC_EMB_WT;
        float v[4], pre[4];
        for (int i = 0; i < 4; ++i) {
            const int col = cg2 * 4 + i;
            pre[i] = c == 1 ? Wt[col] : (c == 2 ? Wt[64 + col] : (c == 3 ? Wt[128 + col] : 0.f));
            v[i] = gelu_erf(pre[i]);
        }
        const float mean = half_warp_sum(v[0] + v[1] + v[2] + v[3]) * (1.0f / 64.0f);
        float var = 0.f;
        for (int i = 0; i < 4; ++i) { v[i] -= mean; var += v[i] * v[i]; }
        var = half_warp_sum(var) * (1.0f / 64.0f);
        const float rs = rsqrtf(var + 1e-5f);
        for (int i = 0; i < 4; ++i) { tA[c * 64 + cg2 * 4 + i] = pre[i]; tXh[c * 64 + cg2 * 4 + i] = v[i] * rs; }
        if (cg2 == 0) tRs[c] = rs;
        lnw[tid] = packed[lay::DEC_EMB + lay::DEC_EMB_LNW + tid];
    }
    __syncthreads();
    Col glw, glb, gw0, gw1, gw2; glw.zero(); glb.zero(); gw0.zero(); gw1.zero(); gw2.zero();
    for (long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long row0 = tile * TM;
        const int n_valid = (int)min((long)TM, n_tok - row0);
        if (tid < TM) {
            const long g = row0 + tid;
            int code = 0;
            if (g < n_tok) code = (g % a.L) == 0 ? 1 : (a.actions[g - 1] != 0.f ? 3 : 2);
            sCode[tid] = code;
        }
        __syncthreads();
        Frag d, xh, h;
        frag_load_g(d, a.gX, row0, n_valid, rg, cg);
        float rsl[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int code = sCode[rg * 4 + i];
            rsl[i] = tRs[code];
#pragma unroll
            for (int j = 0; j < 4; ++j) { xh.v[i][j] = tXh[code * 64 + cg * 4 + j]; h.v[i][j] = tA[code * 64 + cg * 4 + j]; }
        }
        // LN backward with per-row rstd from the table
        col_add_mul(glw, d, xh); col_add(glb, d);
        const float4 w = ld4(lnw, cg);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float gv[4], s1 = 0.f, s2 = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) { gv[j] = d.v[i][j] * f4(w, j); s1 += gv[j]; s2 += gv[j] * xh.v[i][j]; }
            s1 = half_warp_sum(s1) * (1.0f / 64.0f);
            s2 = half_warp_sum(s2) * (1.0f / 64.0f);
            const float rs = (rg * 4 + i) < n_valid ? rsl[i] : 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) d.v[i][j] = rs * (gv[j] - s1 - xh.v[i][j] * s2);
        }
        gelu_bwd(d, h);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int code = sCode[rg * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (code == 1) gw0.v[j] += d.v[i][j];
                else if (code == 2) gw1.v[j] += d.v[i][j];
                else if (code == 3) gw2.v[j] += d.v[i][j];
            }
        }
        __syncthreads();
    }
    col_flush(glw, grad, a.gln_w, cg); col_flush(glb, grad, a.gln_b, cg);
    // action_encoder.0.weight is [64][3] row-major: element (col, k) at col*3 + k
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int col = cg * 4 + j;
        atomicAdd(grad + a.gW + col * 3 + 0, gw0.v[j]);
        atomicAdd(grad + a.gW + col * 3 + 1, gw1.v[j]);
        atomicAdd(grad + a.gW + col * 3 + 2, gw2.v[j]);
    }
}

template <typename K>
int set_smem(K kernel, size_t bytes) {
    CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return MATAS_OK;
}

}  // namespace

// -------------------------------------------------------------------------------------------------
// host orchestration
// -------------------------------------------------------------------------------------------------
int launch_evaluate_backward(const PolicyBufs& b, const TrainBufs& t, const GradMap& gm, float* grad, const float* obs,
                             const float* actions, const float* ava, const float* d_values, const float* d_logp,
                             const float* d_ent, long M, int L, int obs_dim, int mode, int n_sm, cudaStream_t st) {
    if (L > 256) {
        matas_set_error("evaluate backward supports n_agent <= 256 (got %d)", L);
        return MATAS_ERR_LIMIT;
    }
    const size_t bytes = bsmem_bytes(BDEPTH), abytes = bsmem_bytes(ADEPTH);
    // n_agent <= 64: register-tiled attention backward (attn_bwd_tile.cu) + q/k/v input gradients on the tensor cores
    // (bwd_dx_tc.cu); larger n_agent: the half-warp-per-row attention backward inside bwd_attn with SIMT input gradients
    const bool tiled = attn_tile_supported(L);
    float* const* S = t.gS;                              // the current set of layer-output gradient buffers (see take_set)
    auto dx = [&](int kind, int bq, int bk, int bv, int gbq, int gbk, int gbv) -> int {
        DxArgs d;
        d.dq = S[0]; d.dk = S[1]; d.dv = S[2];
        d.blob[0] = bq; d.blob[1] = bk; d.blob[2] = bv;
        d.gb[0] = gbq; d.gb[1] = gbk; d.gb[2] = gbv;
        d.gX = t.gX; d.gRep = t.gRep;
        return launch_bwd_qkv_dx(b.w, grad, d, M * L, kind, n_sm, st);
    };
    const long n_tok = M * L;
    const long n_tiles = (n_tok + TM - 1) / TM;
    const unsigned grid = (unsigned)(n_tiles < n_sm ? n_tiles : n_sm);
    const unsigned agrid = (unsigned)(n_tiles < 2 * n_sm ? n_tiles : 2 * n_sm);     // bwd_attn: two CTAs per SM
    int rc;
    if ((rc = set_smem(bwd_attn<0>, abytes)) || (rc = set_smem(bwd_attn<1>, abytes)) ||
        (rc = set_smem(bwd_head<0>, bytes)) || (rc = set_smem(bwd_head<1>, bytes)) ||
        (rc = set_smem(bwd_embed_enc, bytes)) || (rc = set_smem(bwd_embed_dec, bytes)))
        return rc;
    const float* pk = b.w;
    using namespace lay;
    auto idx = [](int blob) { return (blob - ENC_BLOBS) / BLOB; };
    CUDA_TRY(cudaMemsetAsync(t.gRep, 0, (size_t)n_tok * 64 * sizeof(float), st));
    // Weight gradients: the fused kernels hand their layer-output gradients to
    // bwd_dw_tc.cu through the three scratch buffers t.gS (consumed by the launch that follows each kernel).
    auto job = [](const float* dY, const float* X, int recipe, int ln, int gW) {
        DwJob j; j.dY = dY; j.X = X; j.recipe = recipe; j.ln = ln; j.gW = gW; return j;
    };
    // The weight-gradient launches are DEFERRED: the fused kernels write their layer-output gradients into one of three
    // rotating sets of scratch buffers, the jobs queue up, and one bwd_dw launch takes everything queued when a set is about
    // to be reused (up to two sets = six layers per launch instead of one to three: each launch carries ~35 us of prologue,
    // accumulator read-out and atomics whatever its size).
    DwJobs pending; pending.n = 0;
    int pending_set[8];
    int cur_set = 2;
    auto flush = [&]() -> int {
        if (pending.n == 0) return MATAS_OK;
        const int r = launch_bwd_dw(pk, grad, pending, n_tok, n_sm, st);
        pending.n = 0;
        return r;
    };
    auto take_set = [&]() -> int {                       // the next set becomes S; its queued jobs are launched first
        cur_set = (cur_set + 1) % 3;
        S = t.gS + 3 * cur_set;
        for (int i = 0; i < pending.n; ++i)
            if (pending_set[i] == cur_set) return flush();
        return MATAS_OK;
    };
    auto dw = [&](const DwJob* js, int n) -> int {
        for (int i = 0; i < n; ++i) { pending.j[pending.n] = js[i]; pending_set[pending.n] = cur_set; ++pending.n; }
        return MATAS_OK;
    };

    // ---- decoder ---------------------------------------------------------------------------------
    {
        if ((rc = take_set())) return rc;
        HeadArgs h = {};
        h.d_a = d_logp; h.d_b = d_ent; h.actions = actions; h.ava = ava;
        h.hh = t.d_hh; h.xhh = t.d_xhh; h.rsh = t.d_rsh; h.xh_in = t.d_xh3[1]; h.rs_in = t.d_rs3[1];
        h.ln_in = dec_blob(1, DM2); h.blob_h0 = DEC_H0; h.h3 = DEC_H3;
        h.gW0 = gm.w[idx(DEC_H0)]; h.gb0 = gm.b[idx(DEC_H0)]; h.glnw = gm.lnw[idx(DEC_H0)]; h.glnb = gm.lnb[idx(DEC_H0)];
        h.gW3 = gm.dec_h3_w; h.gb3 = gm.dec_h3_b; h.glogstd = gm.log_std;
        h.gOut = t.gX; h.L = L; h.n_disc = mode == MATAS_MODE_SEMI ? L - 1 : L;
        h.sY = S[0];
        bwd_head<1><<<grid, NT, bytes, st>>>(pk, grad, h, n_tok);
        const DwJob js[1] = {job(S[0], t.d_xh3[1], DW_X_AFFINE, dec_blob(1, DM2), h.gW0)};
        if ((rc = dw(js, 1))) return rc;
    }
    for (int k = 1; k >= 0; --k) {
        if ((rc = take_set())) return rc;
        TailArgs ta = {};
        ta.gOut = t.gX; ta.xhF = t.d_xh3[k]; ta.rsF = t.d_rs3[k]; ta.m0 = t.d_m0[k]; ta.xhM = t.d_xh2[k]; ta.rsM = t.d_rs2[k];
        ta.Y = t.d_y2[k]; ta.lnF = dec_blob(k, DM2); ta.lnM = dec_blob(k, DP2);
        ta.seq[0] = bw(dec_blob(k, DM2)); ta.seq[1] = bw(dec_blob(k, DM0)); ta.seq[2] = bw(dec_blob(k, DP2));
        ta.fW0 = dec_blob(k, DM0);
        ta.gW2 = gm.w[idx(dec_blob(k, DM2))]; ta.gb2 = gm.b[idx(dec_blob(k, DM2))];
        ta.gW0 = gm.w[idx(dec_blob(k, DM0))]; ta.gb0 = gm.b[idx(dec_blob(k, DM0))];
        ta.gWp = gm.w[idx(dec_blob(k, DP2))]; ta.gbp = gm.b[idx(dec_blob(k, DP2))];
        ta.glnFw = gm.lnw[idx(dec_blob(k, DM2))]; ta.glnFb = gm.lnb[idx(dec_blob(k, DM2))];
        ta.glnMw = gm.lnw[idx(dec_blob(k, DP2))]; ta.glnMb = gm.lnb[idx(dec_blob(k, DP2))];
        ta.gRes = t.gRep; ta.res_add = 1; ta.gY = t.gYa; ta.gD = t.gDa; ta.with_mlp = 1;
        ta.sY[0] = S[0]; ta.sY[1] = S[1]; ta.sY[2] = S[2];
        if ((rc = launch_bwd_mlp_tc(pk, grad, ta, n_tok, n_sm, st))) return rc;
        {
            const DwJob js[3] = {job(S[0], t.d_m0[k], DW_X_GELU, 0, ta.gW2), job(S[1], t.d_xh2[k], DW_X_AFFINE, dec_blob(k, DP2), ta.gW0),
                                 job(S[2], t.d_y2[k], DW_X_PLAIN, 0, ta.gWp)};
            if ((rc = dw(js, 3))) return rc;
        }

        if ((rc = take_set())) return rc;
        AttnArgs a2 = {};
        a2.Q = b.Q2[k]; a2.K = b.K2[k]; a2.V = b.V2[k]; a2.lse = t.d_lse2[k]; a2.gY = t.gYa; a2.gD = t.gDa;
        a2.xh = t.d_xh1[k]; a2.rs = t.d_rs1[k]; a2.ln = dec_blob(k, DP1); a2.xin = b.rep;
        a2.seq[0] = bw(dec_q2(k)); a2.seq[1] = bw(dec_blob(k, DK2)); a2.seq[2] = bw(dec_blob(k, DV2));
        a2.gWq = gm.w[idx(dec_q2(k))]; a2.gbq = gm.b[idx(dec_q2(k))];
        a2.gWk = gm.w[idx(dec_blob(k, DK2))]; a2.gbk = gm.b[idx(dec_blob(k, DK2))];
        a2.gWv = gm.w[idx(dec_blob(k, DV2))]; a2.gbv = gm.b[idx(dec_blob(k, DV2))];
        a2.gX = t.gX; a2.gRep = t.gRep; a2.causal = 1; a2.L = L;
        a2.sY[0] = S[0]; a2.sY[1] = S[1]; a2.sY[2] = S[2];
        if (tiled) {
            if ((rc = launch_attn_bwd_tile(a2.Q, a2.K, a2.V, a2.gY, a2.lse, a2.gD, S[0], S[1], S[2], M, L, 1, st))) return rc;
            a2.pre = 1;
        }
        if (tiled) {
            if ((rc = dx(1, dec_q2(k), dec_blob(k, DK2), dec_blob(k, DV2), a2.gbq, a2.gbk, a2.gbv))) return rc;
        } else {
            bwd_attn<1><<<agrid, NT, abytes, st>>>(pk, grad, a2, n_tok);   // gX = grad wrt x1 (LN1 output)
        }
        {
            const DwJob js[3] = {job(S[0], b.rep, DW_X_PLAIN, 0, a2.gWq), job(S[1], t.d_xh1[k], DW_X_AFFINE, dec_blob(k, DP1), a2.gWk),
                                 job(S[2], t.d_xh1[k], DW_X_AFFINE, dec_blob(k, DP1), a2.gWv)};
            if ((rc = dw(js, 3))) return rc;
        }

        if ((rc = take_set())) return rc;
        TailArgs tb = {};
        tb.gOut = t.gX; tb.xhM = t.d_xh1[k]; tb.rsM = t.d_rs1[k]; tb.Y = t.d_y1[k];
        tb.lnF = dec_blob(k, DP1); tb.lnM = dec_blob(k, DP1);
        tb.seq[2] = bw(dec_blob(k, DP1));
        tb.gWp = gm.w[idx(dec_blob(k, DP1))]; tb.gbp = gm.b[idx(dec_blob(k, DP1))];
        tb.glnMw = gm.lnw[idx(dec_blob(k, DP1))]; tb.glnMb = gm.lnb[idx(dec_blob(k, DP1))];
        tb.gRes = t.gX; tb.res_add = 0; tb.gY = t.gYb; tb.gD = t.gDb; tb.with_mlp = 0;
        tb.sY[2] = S[2];
        if ((rc = launch_bwd_mlp_tc(pk, grad, tb, n_tok, n_sm, st))) return rc;
        {
            const DwJob js[1] = {job(S[2], t.d_y1[k], DW_X_PLAIN, 0, tb.gWp)};
            if ((rc = dw(js, 1))) return rc;
        }

        if ((rc = take_set())) return rc;
        AttnArgs a1 = {};
        a1.Q = t.d_q1[k]; a1.K = b.K1[k]; a1.V = b.V1[k]; a1.lse = t.d_lse1[k]; a1.gY = t.gYb; a1.gD = t.gDb;
        if (k == 1) { a1.xh = t.d_xh3[0]; a1.rs = t.d_rs3[0]; a1.ln = dec_blob(0, DM2); }
        else { a1.ln = -1; a1.xin = t.d_x0; }
        a1.seq[0] = bw(dec_blob(k, DQ1)); a1.seq[1] = bw(dec_blob(k, DK1)); a1.seq[2] = bw(dec_blob(k, DV1));
        a1.gWq = gm.w[idx(dec_blob(k, DQ1))]; a1.gbq = gm.b[idx(dec_blob(k, DQ1))];
        a1.gWk = gm.w[idx(dec_blob(k, DK1))]; a1.gbk = gm.b[idx(dec_blob(k, DK1))];
        a1.gWv = gm.w[idx(dec_blob(k, DV1))]; a1.gbv = gm.b[idx(dec_blob(k, DV1))];
        a1.gX = t.gX; a1.causal = 1; a1.L = L;
        a1.sY[0] = S[0]; a1.sY[1] = S[1]; a1.sY[2] = S[2];
        if (tiled) {
            if ((rc = launch_attn_bwd_tile(a1.Q, a1.K, a1.V, a1.gY, a1.lse, a1.gD, S[0], S[1], S[2], M, L, a1.causal, st))) return rc;
            a1.pre = 1;
        }
        if (tiled) {
            if ((rc = dx(0, dec_blob(k, DQ1), dec_blob(k, DK1), dec_blob(k, DV1), a1.gbq, a1.gbk, a1.gbv))) return rc;
        } else {
            bwd_attn<0><<<agrid, NT, abytes, st>>>(pk, grad, a1, n_tok);   // gX += attention path
        }
        {
            const float* X = k == 1 ? t.d_xh3[0] : t.d_x0;
            const int rcp = k == 1 ? DW_X_AFFINE : DW_X_PLAIN, ln = k == 1 ? dec_blob(0, DM2) : 0;
            const DwJob js[3] = {job(S[0], X, rcp, ln, a1.gWq), job(S[1], X, rcp, ln, a1.gWk), job(S[2], X, rcp, ln, a1.gWv)};
            if ((rc = dw(js, 3))) return rc;
        }
    }
    {
        EmbArgs e = {};
        e.gX = t.gX; e.actions = actions; e.L = L;
        e.gln_w = gm.dec_ln_w; e.gln_b = gm.dec_ln_b; e.gW = gm.dec_wa;
        bwd_embed_dec<<<grid, NT, bytes, st>>>(pk, grad, e, n_tok);
    }
    // ---- encoder ---------------------------------------------------------------------------------
    {
        if ((rc = take_set())) return rc;
        HeadArgs h = {};
        h.d_a = d_values; h.hh = t.e_hh; h.xhh = t.e_xhh; h.rsh = t.e_rsh; h.rep = b.rep;
        h.blob_h0 = ENC_H0; h.h3 = ENC_H3;
        h.gW0 = gm.w[idx(ENC_H0)]; h.gb0 = gm.b[idx(ENC_H0)]; h.glnw = gm.lnw[idx(ENC_H0)]; h.glnb = gm.lnb[idx(ENC_H0)];
        h.gW3 = gm.enc_h3_w; h.gb3 = gm.enc_h3_b; h.glogstd = -1;
        h.gOut = t.gRep; h.L = L;
        h.sY = S[0];
        bwd_head<0><<<grid, NT, bytes, st>>>(pk, grad, h, n_tok);
        const DwJob js[1] = {job(S[0], b.rep, DW_X_PLAIN, 0, h.gW0)};
        if ((rc = dw(js, 1))) return rc;
    }
    for (int k = 1; k >= 0; --k) {
        if ((rc = take_set())) return rc;
        TailArgs ta = {};
        ta.gOut = k == 1 ? t.gRep : t.gX; ta.xhF = t.e_xh2[k]; ta.rsF = t.e_rs2[k]; ta.m0 = t.e_m0[k];
        ta.xhM = t.e_xh1[k]; ta.rsM = t.e_rs1[k]; ta.Y = t.e_y[k]; ta.lnF = enc_blob(k, EM2); ta.lnM = enc_blob(k, EP);
        ta.seq[0] = bw(enc_blob(k, EM2)); ta.seq[1] = bw(enc_blob(k, EM0)); ta.seq[2] = bw(enc_blob(k, EP));
        ta.fW0 = enc_blob(k, EM0);
        ta.gW2 = gm.w[idx(enc_blob(k, EM2))]; ta.gb2 = gm.b[idx(enc_blob(k, EM2))];
        ta.gW0 = gm.w[idx(enc_blob(k, EM0))]; ta.gb0 = gm.b[idx(enc_blob(k, EM0))];
        ta.gWp = gm.w[idx(enc_blob(k, EP))]; ta.gbp = gm.b[idx(enc_blob(k, EP))];
        ta.glnFw = gm.lnw[idx(enc_blob(k, EM2))]; ta.glnFb = gm.lnb[idx(enc_blob(k, EM2))];
        ta.glnMw = gm.lnw[idx(enc_blob(k, EP))]; ta.glnMb = gm.lnb[idx(enc_blob(k, EP))];
        ta.gRes = t.gX; ta.res_add = 0; ta.gY = t.gYa; ta.gD = t.gDa; ta.with_mlp = 1;
        ta.sY[0] = S[0]; ta.sY[1] = S[1]; ta.sY[2] = S[2];
        if ((rc = launch_bwd_mlp_tc(pk, grad, ta, n_tok, n_sm, st))) return rc;
        {
            const DwJob js[3] = {job(S[0], t.e_m0[k], DW_X_GELU, 0, ta.gW2), job(S[1], t.e_xh1[k], DW_X_AFFINE, enc_blob(k, EP), ta.gW0),
                                 job(S[2], t.e_y[k], DW_X_PLAIN, 0, ta.gWp)};
            if ((rc = dw(js, 3))) return rc;
        }

        if ((rc = take_set())) return rc;
        AttnArgs a1 = {};
        a1.Q = t.e_q[k]; a1.K = b.K[k]; a1.V = b.V[k]; a1.lse = t.e_lse[k]; a1.gY = t.gYa; a1.gD = t.gDa;
        if (k == 1) { a1.xh = t.e_xh2[0]; a1.rs = t.e_rs2[0]; a1.ln = enc_blob(0, EM2); }
        else { a1.ln = -1; a1.xin = t.e_x0; }
        a1.seq[0] = bw(enc_blob(k, EQ)); a1.seq[1] = bw(enc_blob(k, EK)); a1.seq[2] = bw(enc_blob(k, EV));
        a1.gWq = gm.w[idx(enc_blob(k, EQ))]; a1.gbq = gm.b[idx(enc_blob(k, EQ))];
        a1.gWk = gm.w[idx(enc_blob(k, EK))]; a1.gbk = gm.b[idx(enc_blob(k, EK))];
        a1.gWv = gm.w[idx(enc_blob(k, EV))]; a1.gbv = gm.b[idx(enc_blob(k, EV))];
        a1.gX = t.gX; a1.causal = 0; a1.L = L;
        a1.sY[0] = S[0]; a1.sY[1] = S[1]; a1.sY[2] = S[2];
        if (tiled) {
            if ((rc = launch_attn_bwd_tile(a1.Q, a1.K, a1.V, a1.gY, a1.lse, a1.gD, S[0], S[1], S[2], M, L, a1.causal, st))) return rc;
            a1.pre = 1;
        }
        if (tiled) {
            if ((rc = dx(0, enc_blob(k, EQ), enc_blob(k, EK), enc_blob(k, EV), a1.gbq, a1.gbk, a1.gbv))) return rc;
        } else {
            bwd_attn<0><<<agrid, NT, abytes, st>>>(pk, grad, a1, n_tok);
        }
        {
            const float* X = k == 1 ? t.e_xh2[0] : t.e_x0;
            const int rcp = k == 1 ? DW_X_AFFINE : DW_X_PLAIN, ln = k == 1 ? enc_blob(0, EM2) : 0;
            const DwJob js[3] = {job(S[0], X, rcp, ln, a1.gWq), job(S[1], X, rcp, ln, a1.gWk), job(S[2], X, rcp, ln, a1.gWv)};
            if ((rc = dw(js, 3))) return rc;
        }
    }
    {
        EmbArgs e = {};
        e.gX = t.gX; e.h0 = t.e_h0; e.xh0 = t.e_xh0; e.rs0 = t.e_rs0; e.obs = obs; e.obs_dim = obs_dim;
        e.gln_w = gm.enc_ln_w; e.gln_b = gm.enc_ln_b; e.gW = gm.enc_wobs; e.gb = gm.enc_bobs;
        e.gln7_w = gm.enc_ln7_w; e.gln7_b = gm.enc_ln7_b;
        bwd_embed_enc<<<grid, NT, bytes, st>>>(pk, grad, e, n_tok);
    }
    if ((rc = flush())) return rc;
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}
