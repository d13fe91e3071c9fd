// 64-row tile primitives shared by the policy kernels: the TMA-fed weight ring (every kernel), and the fp32 SIMT
// GEMM stage [64 x 64] * [64 x 64] with fused bias/GELU/residual/LayerNorm epilogue that q2_from_rep and the
// parts of the backward without a tensor-core kernel (heads, embeddings, n_agent > 64 attention) still use.
#pragma once
#include "common.cuh"
#include "layout.h"

namespace tile {
constexpr int TM = 64;        // rows (tokens or envs) per tile
constexpr int XS = 68;        // padded row stride of activation tiles in shared memory (floats)
constexpr int NT = 256;       // threads per CTA: 16 row groups x 16 column groups, 4x4 outputs each
constexpr int TILE_F = TM * XS;

enum { EP_GELU = 1, EP_RES = 2, EP_LN = 4 };

// ---------------------------------------------------------------------------------------------
// weight ring: thread 0 streams the kernel's blob sequence global -> shared with cp.async.bulk,
// completion tracked by one mbarrier per slot; consumers are all NT threads.
// Contract: at most `depth` acquire()s between two prefetch() calls, and a __syncthreads() between the
// last use of a slot and the prefetch() that refills it.
// ---------------------------------------------------------------------------------------------
struct Ring {
    float* slots; uint64_t* bars; const float* gw; const int* seq;   // seq: float offsets of the blobs, cyclic
    int seq_len, total, depth;               // total: blobs this CTA consumes in its lifetime
    uint32_t bytes;                          // bytes per blob (forward blobs: BLOB_BYTES)
    int stride;                              // floats between slots (tensor-core operands need 1024-byte aligned slots)
    int producer;                            // the thread that issues the bulk copies
    // consumer side (every thread keeps the same values): blobs released so far, its slot and that slot's parity.
    // Slot / sequence positions advance by compare-and-wrap: no division by the run-time depth on the stage path.
    int head, head_slot; uint32_t head_par;
    int tail, tail_slot, tail_seq;           // producer side: blobs requested so far, the next slot / sequence position

    __device__ void init(float* slots_, uint64_t* bars_, const float* gw_, const int* seq_, int seq_len_, int total_,
                         int depth_, uint32_t bytes_ = lay::BLOB_BYTES, int stride_ = lay::BLOB, int producer_ = 0) {
        slots = slots_; bars = bars_; gw = gw_; seq = seq_; seq_len = seq_len_; total = total_; depth = depth_;
        bytes = bytes_; stride = stride_; producer = producer_;
        head = 0; head_slot = 0; head_par = 0; tail = 0; tail_slot = 0; tail_seq = 0;
        if (threadIdx.x == 0) {
            for (int i = 0; i < depth; ++i) mbar_init(&bars[i], 1);
            mbar_fence_init();
        }
    }
    // Refill every slot released so far.  A slot is rewritten only after the producer itself has passed the stage that
    // consumed it (it waits for the MMAs like every thread) and a CTA barrier has closed that stage's epilogue reads.
    __device__ void prefetch() {
        if ((int)threadIdx.x == producer) {
            while (tail < head + depth && tail < total) {
                mbar_expect_tx(&bars[tail_slot], bytes);
                bulk_g2s(slots + (size_t)tail_slot * stride, gw + seq[tail_seq], bytes, &bars[tail_slot]);
                ++tail;
                if (++tail_slot == depth) tail_slot = 0;
                if (++tail_seq == seq_len) tail_seq = 0;
            }
        }
    }
    __device__ const float* acquire() {
        mbar_wait(&bars[head_slot], head_par);
        return slots + (size_t)head_slot * stride;
    }
    // k-th blob after the current head without consuming it (stages that need two blobs at once); k < depth
    __device__ const float* peek(int k) {
        int s = head_slot + k;
        uint32_t par = head_par;
        if (s >= depth) { s -= depth; par ^= 1u; }
        mbar_wait(&bars[s], par);
        return slots + (size_t)s * stride;
    }
    __device__ void release() {
        ++head;
        if (++head_slot == depth) { head_slot = 0; head_par ^= 1u; }
    }
};

// ---------------------------------------------------------------------------------------------
// GEMM: acc[i][j] = sum_k X[RPT*rg+i][k] * Wt[k][4*cg+j]     (tile of 16*RPT rows; RPT rows per thread)
// ---------------------------------------------------------------------------------------------
template <int RPT>
__device__ __forceinline__ void gemm(const float* __restrict__ X, const float* __restrict__ Wt, float (&acc)[RPT][4],
                                     int rg, int cg) {
#pragma unroll
    for (int i = 0; i < RPT; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    const float* xr = X + rg * RPT * XS;
    const float* wc = Wt + cg * 4;
#pragma unroll 4
    for (int k = 0; k < 64; k += 4) {
        float4 a[RPT], w[4];
#pragma unroll
        for (int i = 0; i < RPT; ++i) a[i] = *reinterpret_cast<const float4*>(xr + i * XS + k);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) w[kk] = *reinterpret_cast<const float4*>(wc + (k + kk) * 64);
#pragma unroll
        for (int i = 0; i < RPT; ++i) {
            const float av[4] = {a[i].x, a[i].y, a[i].z, a[i].w};
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                acc[i][0] = fmaf(av[kk], w[kk].x, acc[i][0]);
                acc[i][1] = fmaf(av[kk], w[kk].y, acc[i][1]);
                acc[i][2] = fmaf(av[kk], w[kk].z, acc[i][2]);
                acc[i][3] = fmaf(av[kk], w[kk].w, acc[i][3]);
            }
        }
    }
}

// Optional training-mode outputs of a stage (global, token-major; null = not saved):
//   pre  [row][64]  bias-added pre-activation (input of GELU)
//   xhat [row][64]  LayerNorm-normalised value before the affine map;  rstd [row]
struct SaveOut { float* pre; float* xhat; float* rstd; };

// Epilogue order: + bias -> GELU -> + residual -> LayerNorm (eps 1e-5, biased variance).
// The residual tile (shared, stride XS, or global) is read into registers BEFORE the GEMM so that
// out may alias it and global latency hides behind the FMAs; rows >= res_valid read as 0.
// Output row r goes to out + r*out_stride (shared or global); rows >= n_valid are not stored.
template <int FLAGS, int RPT = 4>
__device__ __forceinline__ void stage(Ring& ring, const float* __restrict__ X, const float* res, long res_stride,
                                      int res_valid, float* out, long out_stride, int n_valid, int rg, int cg,
                                      const SaveOut* sv = nullptr, int sv_valid = 0) {
    float4 rr[RPT];
    if (FLAGS & EP_RES) {
#pragma unroll
        for (int i = 0; i < RPT; ++i) {
            const int r = rg * RPT + i;
            rr[i] = r < res_valid ? *reinterpret_cast<const float4*>(res + (long)r * res_stride + cg * 4)
                                  : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    const float* blob = ring.acquire();
    float acc[RPT][4];
    gemm<RPT>(X, blob, acc, rg, cg);
    const float4 b = *reinterpret_cast<const float4*>(blob + lay::B_BIAS + cg * 4);
    float4 lw, lb;
    if (FLAGS & EP_LN) {
        lw = *reinterpret_cast<const float4*>(blob + lay::B_LNW + cg * 4);
        lb = *reinterpret_cast<const float4*>(blob + lay::B_LNB + cg * 4);
    }
    ring.release();
#pragma unroll
    for (int i = 0; i < RPT; ++i) {
        const int r = rg * RPT + i;
        float v0 = acc[i][0] + b.x, v1 = acc[i][1] + b.y, v2 = acc[i][2] + b.z, v3 = acc[i][3] + b.w;
        if ((FLAGS & EP_GELU) && sv && sv->pre && r < sv_valid)
            *reinterpret_cast<float4*>(sv->pre + (long)r * 64 + cg * 4) = make_float4(v0, v1, v2, v3);
        if (FLAGS & EP_GELU) { v0 = gelu_erf(v0); v1 = gelu_erf(v1); v2 = gelu_erf(v2); v3 = gelu_erf(v3); }
        if (FLAGS & EP_RES) { v0 += rr[i].x; v1 += rr[i].y; v2 += rr[i].z; v3 += rr[i].w; }
        if (FLAGS & EP_LN) {
            const float mean = half_warp_sum(v0 + v1 + v2 + v3) * (1.0f / 64.0f);
            const float d0 = v0 - mean, d1 = v1 - mean, d2 = v2 - mean, d3 = v3 - mean;
            const float var = half_warp_sum(d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3) * (1.0f / 64.0f);
            const float rs = rsqrtf(var + 1e-5f);
            if (sv && sv->xhat && r < sv_valid) {
                *reinterpret_cast<float4*>(sv->xhat + (long)r * 64 + cg * 4) = make_float4(d0 * rs, d1 * rs, d2 * rs, d3 * rs);
                if (cg == 0) sv->rstd[r] = rs;
            }
            v0 = d0 * rs * lw.x + lb.x; v1 = d1 * rs * lw.y + lb.y; v2 = d2 * rs * lw.z + lb.z; v3 = d3 * rs * lw.w + lb.w;
        }
        if (r < n_valid) *reinterpret_cast<float4*>(out + (long)r * out_stride + cg * 4) = make_float4(v0, v1, v2, v3);
    }
}
// plain stage (no residual)
template <int FLAGS, int RPT = 4>
__device__ __forceinline__ void stage(Ring& ring, const float* __restrict__ X, float* out, long out_stride, int n_valid,
                                      int rg, int cg, const SaveOut* sv = nullptr, int sv_valid = 0) {
    stage<FLAGS, RPT>(ring, X, nullptr, 0, 0, out, out_stride, n_valid, rg, cg, sv, sv_valid);
}

// load a [16*RPT][64] tile from global rows (row r at src + r*stride) into a shared tile; invalid rows -> 0
template <int RPT = 4>
__device__ __forceinline__ void load_tile(float* __restrict__ dst, const float* __restrict__ src, long stride, int n_valid) {
    constexpr int N = 16 * RPT * 16 / NT;             // all loads in flight before the first shared store
    float4 v[N];
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const int idx = threadIdx.x + i * NT, r = idx >> 4, c = idx & 15;
        v[i] = r < n_valid ? __ldcg(reinterpret_cast<const float4*>(src + (long)r * stride + c * 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const int idx = threadIdx.x + i * NT, r = idx >> 4, c = idx & 15;
        *reinterpret_cast<float4*>(dst + r * XS + c * 4) = v[i];
    }
}

// out[r] = dot(X[r][0:64], w) + b for r < n_valid (one thread per row)
__device__ __forceinline__ float row_dot(const float* __restrict__ X, int r, const float* __restrict__ w) {
    const float4* xp = reinterpret_cast<const float4*>(X + r * XS);
    const float4* wp = reinterpret_cast<const float4*>(w);
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const float4 x = xp[i], ww = wp[i];
        s = fmaf(x.x, ww.x, s); s = fmaf(x.y, ww.y, s); s = fmaf(x.z, ww.z, s); s = fmaf(x.w, ww.w, s);
    }
    return s;
}
// -------------------------------------------------------------------------------------------------
// action-token embedding table: emb[c] = LN(GELU(Wa . vec(c))) (ma_transformer.py:224-225)
//   code 0: zero row, 1: start token (1,0,0), 2/3: one-hot of action 0/1, 4/5: start + one-hot
// written to dst[6][64] by the first 96 threads (one half-warp per code)
// -------------------------------------------------------------------------------------------------
__device__ __forceinline__ void build_emb_table(float* __restrict__ dst, const float* __restrict__ w) {
    const int tid = threadIdx.x;
    if (tid < 96) {
        const int c = tid >> 4, cg = tid & 15;
        const float f0 = (c == 1 || c >= 4) ? 1.f : 0.f, f1 = (c == 2 || c == 4) ? 1.f : 0.f, f2 = (c == 3 || c == 5) ? 1.f : 0.f;
        const float* Wt = w + lay::DEC_EMB + lay::DEC_EMB_WT;
        float v[4];
        for (int i = 0; i < 4; ++i) {
            const int col = cg * 4 + i;
            v[i] = gelu_erf(f0 * Wt[col] + f1 * Wt[64 + col] + f2 * Wt[128 + col]);
        }
        const float mean = half_warp_sum(v[0] + v[1] + v[2] + v[3]) * (1.0f / 64.0f);
        float var = 0.f;
        for (int i = 0; i < 4; ++i) { v[i] -= mean; var += v[i] * v[i]; }
        var = half_warp_sum(var) * (1.0f / 64.0f);
        const float rs = rsqrtf(var + 1e-5f);
        for (int i = 0; i < 4; ++i) {
            const int col = cg * 4 + i;
            dst[c * 64 + col] = v[i] * rs * w[lay::DEC_EMB + lay::DEC_EMB_LNW + col] + w[lay::DEC_EMB + lay::DEC_EMB_LNB + col];
        }
    }
}
}  // namespace tile
