// 64-row tile primitives shared by every policy kernel: TMA-fed weight ring, fp32 SIMT GEMM
// [64 x 64] * [64 x 64] with fused bias/GELU/residual/LayerNorm epilogue, and row-wise attention
// over K/V that live in global memory (L2-resident scratch).
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
    int seq_len, total, head, tail, depth;   // total: blobs this CTA consumes in its lifetime; tail: thread 0 only
    uint32_t bytes;                          // bytes per blob (forward blobs: BLOB_BYTES)
    int stride;                              // floats between slots (tensor-core operands need 1024-byte aligned slots)

    __device__ void init(float* slots_, uint64_t* bars_, const float* gw_, const int* seq_, int seq_len_, int total_,
                         int depth_, uint32_t bytes_ = lay::BLOB_BYTES, int stride_ = lay::BLOB) {
        slots = slots_; bars = bars_; gw = gw_; seq = seq_; seq_len = seq_len_; total = total_; depth = depth_;
        head = 0; tail = 0; bytes = bytes_; stride = stride_;
        if (threadIdx.x == 0) {
            for (int i = 0; i < depth; ++i) mbar_init(&bars[i], 1);
            mbar_fence_init();
        }
    }
    __device__ void prefetch() {
        if (threadIdx.x == 0) {
            while (tail < head + depth && tail < total) {
                const int s = tail % depth;
                mbar_expect_tx(&bars[s], bytes);
                bulk_g2s(slots + (size_t)s * stride, gw + seq[tail % seq_len], bytes, &bars[s]);
                ++tail;
            }
        }
    }
    __device__ const float* acquire() {
        const int s = head % depth;
        mbar_wait(&bars[s], (head / depth) & 1);
        return slots + (size_t)s * stride;
    }
    // k-th blob after the current head without consuming it (stages that need two blobs at once)
    __device__ const float* peek(int k) {
        const int h = head + k, s = h % depth;
        mbar_wait(&bars[s], (h / depth) & 1);
        return slots + (size_t)s * stride;
    }
    __device__ void release() { ++head; }
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
// Two-input stage: acc = X0 . W(blob n) + X1 . W(blob n+1); bias / LayerNorm come from the second blob.
template <int FLAGS, int RPT = 4>
__device__ __forceinline__ void stage2(Ring& ring, const float* __restrict__ X0, const float* __restrict__ X1,
                                       const float* res, long res_stride, int res_valid, float* out, long out_stride,
                                       int n_valid, int rg, int cg) {
    float4 rr[RPT];
    if (FLAGS & EP_RES) {
#pragma unroll
        for (int i = 0; i < RPT; ++i) {
            const int r = rg * RPT + i;
            rr[i] = r < res_valid ? *reinterpret_cast<const float4*>(res + (long)r * res_stride + cg * 4)
                                  : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    float acc[RPT][4], acc1[RPT][4];
    const float* blob0 = ring.acquire();
    gemm<RPT>(X0, blob0, acc, rg, cg);
    ring.release();
    const float* blob = ring.acquire();
    gemm<RPT>(X1, blob, acc1, rg, cg);
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
        float v0 = acc[i][0] + acc1[i][0] + b.x, v1 = acc[i][1] + acc1[i][1] + b.y;
        float v2 = acc[i][2] + acc1[i][2] + b.z, v3 = acc[i][3] + acc1[i][3] + b.w;
        if (FLAGS & EP_RES) { v0 += rr[i].x; v1 += rr[i].y; v2 += rr[i].z; v3 += rr[i].w; }
        if (FLAGS & EP_LN) {
            const float mean = half_warp_sum(v0 + v1 + v2 + v3) * (1.0f / 64.0f);
            const float d0 = v0 - mean, d1 = v1 - mean, d2 = v2 - mean, d3 = v3 - mean;
            const float var = half_warp_sum(d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3) * (1.0f / 64.0f);
            const float rs = rsqrtf(var + 1e-5f);
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

// ---------------------------------------------------------------------------------------------
// attention over rows that live in global memory (K/V of the token-parallel passes, cached layer
// inputs of the decode); head h uses columns 32h..32h+31 of a 64-float row.
// ---------------------------------------------------------------------------------------------
// K/V loads.  CACHED (token-parallel passes): the rows were written by an EARLIER kernel, so the
// read-only path (ld.global.nc) is legal and the rows of an env are shared by all query rows of a tile.
// !CACHED (decode): the rows were written earlier in THIS kernel by this CTA, so a plain coherent load
// is required; it may still hit in L1 (same-SM stores keep L1 coherent), which merges the eight
// 16-byte requests a thread makes to one 128-byte row.
template <bool CACHED> __device__ __forceinline__ float4 ld_kv(const float4* p) { return CACHED ? __ldg(p) : *p; }

// Attention for a tile of query rows over K/V rows in global memory (token-parallel passes).
// One half-warp per query row: lane l holds columns 4l..4l+3 of the row (lanes 0-7 = head 0, lanes
// 8-15 = head 1), so every K / V row is read coalesced (one float4 per lane, 256 B per half-warp) and
// exactly once; the 8 lanes of a head reduce their partial dots with three shuffles and keep an
// online softmax (running max, sum, context chunk) in registers.  The key loop is software-pipelined
// (next 4 keys in flight).  No score buffer and no CTA barrier until sOut is published.
// Row r attends to keys [0, nt[r]) at kptr[r] / vptr[r] (t-major, 64 floats per key).
template <int RPT = 4, bool CACHED = false>
__device__ __forceinline__ void attention_rows(const float* __restrict__ sQ, float* __restrict__ sOut,
                                               const float* const* __restrict__ kptr, const float* const* __restrict__ vptr,
                                               const int* __restrict__ nt, float* lse_out = nullptr, int lse_valid = 0) {
    constexpr int TM = 16 * RPT;
    const float scale = 0.17677669529663687f * 1.4426950408889634f;   // log2(e)/sqrt(32): softmax in the exp2 domain
    const int tid = threadIdx.x, hw = tid >> 4, l = tid & 15;
    const unsigned hmask = 0xffffu << (tid & 16);
    for (int r = hw; r < TM; r += NT / 16) {
        const int nr = nt[r];
        float4 q = *reinterpret_cast<const float4*>(sQ + r * XS + l * 4);
        q.x *= scale; q.y *= scale; q.z *= scale; q.w *= scale;
        const float4* kp = reinterpret_cast<const float4*>(kptr[r]) + l;      // key t at kp[t * 16]
        const float4* vp = reinterpret_cast<const float4*>(vptr[r]) + l;
        float m = -INFINITY, sum = 0.f;
        float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
        auto batch = [&](const float4 (&k)[4], const float4 (&v)[4], int t0) {
            float s[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float p = q.x * k[i].x + q.y * k[i].y + q.z * k[i].z + q.w * k[i].w;
                p += __shfl_xor_sync(hmask, p, 4);
                p += __shfl_xor_sync(hmask, p, 2);
                p += __shfl_xor_sync(hmask, p, 1);
                s[i] = t0 + i < nr ? p : -INFINITY;
            }
            const float n = fmaxf(m, fmaxf(fmaxf(s[0], s[1]), fmaxf(s[2], s[3])));
            const float a = exp2f(m - n);                   // 2^(-inf) = 0 on the first batch
            sum *= a; c.x *= a; c.y *= a; c.z *= a; c.w *= a;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float p = exp2f(s[i] - n);
                sum += p;
                c.x = fmaf(p, v[i].x, c.x); c.y = fmaf(p, v[i].y, c.y); c.z = fmaf(p, v[i].z, c.z); c.w = fmaf(p, v[i].w, c.w);
            }
            m = n;
        };
        auto fetch = [&](float4 (&k)[4], float4 (&v)[4], int t0) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const bool ok = t0 + i < nr;
                k[i] = ok ? ld_kv<CACHED>(kp + (long)(t0 + i) * 16) : make_float4(0.f, 0.f, 0.f, 0.f);
                v[i] = ok ? ld_kv<CACHED>(vp + (long)(t0 + i) * 16) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        };
        float4 ka[4], va[4], kb[4], vb[4];
        fetch(ka, va, 0);
        for (int t0 = 0; t0 < nr; t0 += 8) {
            fetch(kb, vb, t0 + 4);
            batch(ka, va, t0);
            fetch(ka, va, t0 + 8);
            if (t0 + 4 < nr) batch(kb, vb, t0 + 4);
        }
        const float inv = nr > 0 ? 1.0f / sum : 0.f;
        *reinterpret_cast<float4*>(sOut + r * XS + l * 4) = make_float4(c.x * inv, c.y * inv, c.z * inv, c.w * inv);
        // natural-log log-sum-exp of the scaled scores (m is in log2 units), training only
        if ((l & 7) == 0 && lse_out && r < lse_valid) lse_out[r * 2 + (l >> 3)] = m * 0.6931471805599453f + logf(sum);
    }
    __syncthreads();
}

// L2 prefetch of the first n_rows cache rows (256 B each) of every valid env slot of the tile: issued
// ahead of the GEMM stages that precede an attention, so its loads find the rows in L2 instead of HBM.
template <int RPT>
__device__ __forceinline__ void prefetch_cache_rows(const float* base, long slot_stride, int n_valid, int n_rows) {
    const int lines = n_rows * 2;       // 128-byte lines per env
    for (int idx = threadIdx.x; idx < n_valid * lines; idx += NT) {
        const int r = idx / lines, l = idx - r * lines;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(base + (long)r * slot_stride + l * 32));
    }
}

// Absorbed attention (decode).  The cache row t of query row r is the 64-float INPUT x_t of this
// attention (xptr[r] + t*64, written earlier by this CTA); sQ0 / sQ1 hold the folded queries q'_h
// (log2(e)/sqrt(32) and Wk_h^T already applied: exp2-domain scores), so score_h(t) = q'_h . x_t and
// ctx_h = sum_t p_h(t) x_t; the
// folded output maps turn (ctx_0, ctx_1) into the attention block's output.  One cache row serves
// both heads.  Outputs overwrite sQ0 / sQ1 with ctx_0 / ctx_1.
template <int RPT = 4>
__device__ __forceinline__ void attention_absorbed(float* __restrict__ sQ0, float* __restrict__ sQ1,
                                                   const float* const* __restrict__ xptr, const int* __restrict__ nt) {
    // Eight lanes per env row: lane g holds columns 4g..4g+3 and 32+4g..32+4g+3 of every 64-float row, so
    // the 8 lanes read each 256-byte cache row as two fully used 128-byte lines (coalesced, ONE pass over
    // the cache), reduce the two heads' partial dots with 4 shuffles, and keep an online softmax (running
    // max / sum / context chunks) in registers.  Fewer lanes per row than a half-warp halves the
    // replicated per-lane work (shuffles, max, exp2, rescale), which is what bounds this loop (issue slots).
    // The key loop is software-pipelined.  No score buffer, no CTA barrier until the contexts are published.
    constexpr int TM = 16 * RPT;
    constexpr int KB = 2;                     // keys per batch (register budget: 128 per thread at 2 CTAs/SM)
    const int tid = threadIdx.x, grp = tid >> 3, g = tid & 7;
    const unsigned gmask = 0xffu << (tid & 24);
    const bool up = (g & 4) != 0;
    for (int r = grp; r < TM; r += NT / 8) {
        const int nr = nt[r];
        const float4 qa0 = *reinterpret_cast<const float4*>(sQ0 + r * XS + g * 4);
        const float4 qb0 = *reinterpret_cast<const float4*>(sQ0 + r * XS + 32 + g * 4);
        const float4 qa1 = *reinterpret_cast<const float4*>(sQ1 + r * XS + g * 4);
        const float4 qb1 = *reinterpret_cast<const float4*>(sQ1 + r * XS + 32 + g * 4);
        const float4* xp = reinterpret_cast<const float4*>(xptr[r]) + g;      // row t: xp[t*16] and xp[t*16 + 8]
        float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
        float4 ca0 = make_float4(0.f, 0.f, 0.f, 0.f), cb0 = ca0, ca1 = ca0, cb1 = ca0;
        auto batch = [&](const float4 (&xa)[KB], const float4 (&xb)[KB], int t0) {
            float s0[KB], s1[KB];
#pragma unroll
            for (int i = 0; i < KB; ++i) {
                const float p0 = qa0.x * xa[i].x + qa0.y * xa[i].y + qa0.z * xa[i].z + qa0.w * xa[i].w +
                                 qb0.x * xb[i].x + qb0.y * xb[i].y + qb0.z * xb[i].z + qb0.w * xb[i].w;
                const float p1 = qa1.x * xa[i].x + qa1.y * xa[i].y + qa1.z * xa[i].z + qa1.w * xa[i].w +
                                 qb1.x * xb[i].x + qb1.y * xb[i].y + qb1.z * xb[i].z + qb1.w * xb[i].w;
                // packed reduction: lanes 0-3 collect head 0, lanes 4-7 head 1, then swap the totals
                float v = up ? p1 : p0;
                v += __shfl_xor_sync(gmask, up ? p0 : p1, 4);
                v += __shfl_xor_sync(gmask, v, 2);
                v += __shfl_xor_sync(gmask, v, 1);
                const float o = __shfl_xor_sync(gmask, v, 4);
                const bool ok = t0 + i < nr;
                s0[i] = ok ? (up ? o : v) : -INFINITY;
                s1[i] = ok ? (up ? v : o) : -INFINITY;
            }
            float n0 = m0, n1 = m1;
#pragma unroll
            for (int i = 0; i < KB; ++i) { n0 = fmaxf(n0, s0[i]); n1 = fmaxf(n1, s1[i]); }
            const float a0 = exp2f(m0 - n0), a1 = exp2f(m1 - n1);      // 2^(-inf) = 0 on the first batch
            l0 *= a0; l1 *= a1;
            ca0.x *= a0; ca0.y *= a0; ca0.z *= a0; ca0.w *= a0; cb0.x *= a0; cb0.y *= a0; cb0.z *= a0; cb0.w *= a0;
            ca1.x *= a1; ca1.y *= a1; ca1.z *= a1; ca1.w *= a1; cb1.x *= a1; cb1.y *= a1; cb1.z *= a1; cb1.w *= a1;
#pragma unroll
            for (int i = 0; i < KB; ++i) {
                const float p0 = exp2f(s0[i] - n0), p1 = exp2f(s1[i] - n1);
                l0 += p0; l1 += p1;
                ca0.x = fmaf(p0, xa[i].x, ca0.x); ca0.y = fmaf(p0, xa[i].y, ca0.y); ca0.z = fmaf(p0, xa[i].z, ca0.z); ca0.w = fmaf(p0, xa[i].w, ca0.w);
                cb0.x = fmaf(p0, xb[i].x, cb0.x); cb0.y = fmaf(p0, xb[i].y, cb0.y); cb0.z = fmaf(p0, xb[i].z, cb0.z); cb0.w = fmaf(p0, xb[i].w, cb0.w);
                ca1.x = fmaf(p1, xa[i].x, ca1.x); ca1.y = fmaf(p1, xa[i].y, ca1.y); ca1.z = fmaf(p1, xa[i].z, ca1.z); ca1.w = fmaf(p1, xa[i].w, ca1.w);
                cb1.x = fmaf(p1, xb[i].x, cb1.x); cb1.y = fmaf(p1, xb[i].y, cb1.y); cb1.z = fmaf(p1, xb[i].z, cb1.z); cb1.w = fmaf(p1, xb[i].w, cb1.w);
            }
            m0 = n0; m1 = n1;
        };
        auto fetch = [&](float4 (&xa)[KB], float4 (&xb)[KB], int t0) {
#pragma unroll
            for (int i = 0; i < KB; ++i) {
                const bool ok = t0 + i < nr;
                xa[i] = ok ? xp[(long)(t0 + i) * 16] : make_float4(0.f, 0.f, 0.f, 0.f);
                xb[i] = ok ? xp[(long)(t0 + i) * 16 + 8] : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        };
        // software pipeline: the next KB rows are in flight while the current KB are reduced
        float4 xa[KB], xb[KB], ya[KB], yb[KB];
        fetch(xa, xb, 0);
        for (int t0 = 0; t0 < nr; t0 += 2 * KB) {
            fetch(ya, yb, t0 + KB);
            batch(xa, xb, t0);
            fetch(xa, xb, t0 + 2 * KB);
            if (t0 + KB < nr) batch(ya, yb, t0 + KB);
        }
        const float i0 = nr > 0 ? 1.0f / l0 : 0.f, i1 = nr > 0 ? 1.0f / l1 : 0.f;
        *reinterpret_cast<float4*>(sQ0 + r * XS + g * 4) = make_float4(ca0.x * i0, ca0.y * i0, ca0.z * i0, ca0.w * i0);
        *reinterpret_cast<float4*>(sQ0 + r * XS + 32 + g * 4) = make_float4(cb0.x * i0, cb0.y * i0, cb0.z * i0, cb0.w * i0);
        *reinterpret_cast<float4*>(sQ1 + r * XS + g * 4) = make_float4(ca1.x * i1, ca1.y * i1, ca1.z * i1, ca1.w * i1);
        *reinterpret_cast<float4*>(sQ1 + r * XS + 32 + g * 4) = make_float4(cb1.x * i1, cb1.y * i1, cb1.z * i1, cb1.w * i1);
    }
    __syncthreads();
}

// load a [16*RPT][64] tile from global rows (row r at src + r*stride) into a shared tile; invalid rows -> 0
template <int RPT = 4>
__device__ __forceinline__ void load_tile(float* __restrict__ dst, const float* __restrict__ src, long stride, int n_valid) {
    for (int idx = threadIdx.x; idx < 16 * RPT * 16; idx += NT) {
        const int r = idx >> 4, c = idx & 15;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < n_valid) v = __ldcg(reinterpret_cast<const float4*>(src + (long)r * stride + c * 4));
        *reinterpret_cast<float4*>(dst + r * XS + c * 4) = v;
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
