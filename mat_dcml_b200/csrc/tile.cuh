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

    __device__ void init(float* slots_, uint64_t* bars_, const float* gw_, const int* seq_, int seq_len_, int total_,
                         int depth_, uint32_t bytes_ = lay::BLOB_BYTES) {
        slots = slots_; bars = bars_; gw = gw_; seq = seq_; seq_len = seq_len_; total = total_; depth = depth_;
        head = 0; tail = 0; bytes = bytes_;
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
                bulk_g2s(slots + (size_t)s * lay::BLOB, gw + seq[tail % seq_len], bytes, &bars[s]);
                ++tail;
            }
        }
    }
    __device__ const float* acquire() {
        const int s = head % depth;
        mbar_wait(&bars[s], (head / depth) & 1);
        return slots + (size_t)s * lay::BLOB;
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
// attention for a tile of query rows.  Row r attends to nt[r] keys K[r][t][:], V[r][t][:]
// (t-major, 64 floats per key: head h uses columns 32h..32h+31) found at kptr[r] / vptr[r].
// sScore: TM*2*lpad floats (lpad odd), sInv: TM*2 floats.  Three phases with CTA barriers.
// ---------------------------------------------------------------------------------------------
// K/V loads.  CACHED (token-parallel passes): the rows were written by an EARLIER kernel, so the
// read-only path (ld.global.nc) is legal and the rows of an env are shared by all query rows of a tile.
// !CACHED (decode): the rows were written earlier in THIS kernel by this CTA, so a plain coherent load
// is required; it may still hit in L1 (same-SM stores keep L1 coherent), which merges the eight
// 16-byte requests a thread makes to one 128-byte row.
template <bool CACHED> __device__ __forceinline__ float4 ld_kv(const float4* p) { return CACHED ? __ldg(p) : *p; }

__device__ __forceinline__ float dot32(const float4 (&q)[8], const float4 (&k)[8]) {
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;     // four independent chains, 8 deep each
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        s0 = fmaf(q[i].x, k[i].x, s0); s1 = fmaf(q[i].y, k[i].y, s1);
        s2 = fmaf(q[i].z, k[i].z, s2); s3 = fmaf(q[i].w, k[i].w, s3);
    }
    return (s0 + s1) + (s2 + s3);
}

// One (row, head) pair is owned by G = NT / (2 TM) adjacent lanes, which split its keys round-robin:
// scores (q in registers, two key rows in flight) -> max / exp / sum with shuffles over the G lanes
// -> P.V accumulated in registers and all-reduced over the G lanes.  No CTA barrier inside; one at
// the end publishes sOut.  sScore holds the pair's scores (each lane only re-reads its own entries).
template <int RPT = 4, bool CACHED = false>
__device__ __forceinline__ void attention(const float* __restrict__ sQ, float* __restrict__ sOut,
                                          float* __restrict__ sScore, float* __restrict__ sInv, int lpad,
                                          const float* const* __restrict__ kptr, const float* const* __restrict__ vptr,
                                          const int* __restrict__ nt, int nt_max, float* lse_out = nullptr,
                                          int lse_valid = 0) {
    constexpr int TM = 16 * RPT;
    constexpr int G = NT / (TM * 2);
    const float scale = 0.17677669529663687f;   // 1/sqrt(32)
    const int tid = threadIdx.x, rh = tid / G, sub = tid % G, r = rh >> 1, h = rh & 1;
    const int n = nt[r];
    const float4* kp = reinterpret_cast<const float4*>(kptr[r] + h * 32);
    const float4* vp = reinterpret_cast<const float4*>(vptr[r] + h * 32);
    float* sc = sScore + rh * lpad;
    float m = -INFINITY;
    {
        float4 q[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) q[i] = *reinterpret_cast<const float4*>(sQ + r * XS + h * 32 + i * 4);
        int t = sub;
        for (; t + G < n; t += 2 * G) {
            float4 k0[8], k1[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) k0[i] = ld_kv<CACHED>(kp + (long)t * 16 + i);
#pragma unroll
            for (int i = 0; i < 8; ++i) k1[i] = ld_kv<CACHED>(kp + (long)(t + G) * 16 + i);
            const float a0 = dot32(q, k0) * scale, a1 = dot32(q, k1) * scale;
            sc[t] = a0; sc[t + G] = a1;
            m = fmaxf(m, fmaxf(a0, a1));
        }
        if (t < n) {
            float4 k0[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) k0[i] = ld_kv<CACHED>(kp + (long)t * 16 + i);
            const float a0 = dot32(q, k0) * scale;
            sc[t] = a0;
            m = fmaxf(m, a0);
        }
    }
#pragma unroll
    for (int o = 1; o < G; o <<= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float4 acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    float sum = 0.f;
    {
        int t = sub;
        for (; t + G < n; t += 2 * G) {
            float4 v0[8], v1[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v0[i] = ld_kv<CACHED>(vp + (long)t * 16 + i);
#pragma unroll
            for (int i = 0; i < 8; ++i) v1[i] = ld_kv<CACHED>(vp + (long)(t + G) * 16 + i);
            const float p0 = expf(sc[t] - m), p1 = expf(sc[t + G] - m);
            sum += p0 + p1;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                acc[i].x = fmaf(p0, v0[i].x, acc[i].x); acc[i].y = fmaf(p0, v0[i].y, acc[i].y);
                acc[i].z = fmaf(p0, v0[i].z, acc[i].z); acc[i].w = fmaf(p0, v0[i].w, acc[i].w);
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                acc[i].x = fmaf(p1, v1[i].x, acc[i].x); acc[i].y = fmaf(p1, v1[i].y, acc[i].y);
                acc[i].z = fmaf(p1, v1[i].z, acc[i].z); acc[i].w = fmaf(p1, v1[i].w, acc[i].w);
            }
        }
        if (t < n) {
            float4 v0[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v0[i] = ld_kv<CACHED>(vp + (long)t * 16 + i);
            const float p0 = expf(sc[t] - m);
            sum += p0;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                acc[i].x = fmaf(p0, v0[i].x, acc[i].x); acc[i].y = fmaf(p0, v0[i].y, acc[i].y);
                acc[i].z = fmaf(p0, v0[i].z, acc[i].z); acc[i].w = fmaf(p0, v0[i].w, acc[i].w);
            }
        }
    }
#pragma unroll
    for (int o = 1; o < G; o <<= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            acc[i].x += __shfl_xor_sync(0xffffffffu, acc[i].x, o); acc[i].y += __shfl_xor_sync(0xffffffffu, acc[i].y, o);
            acc[i].z += __shfl_xor_sync(0xffffffffu, acc[i].z, o); acc[i].w += __shfl_xor_sync(0xffffffffu, acc[i].w, o);
        }
    }
    const float inv = n > 0 ? 1.0f / sum : 0.f;
    // every lane holds the totals; lane `sub` stores its 8/G float4 share of the 32 outputs
#pragma unroll
    for (int i = 0; i < 8; ++i)
        if (i / (8 / G) == sub)
            *reinterpret_cast<float4*>(sOut + r * XS + h * 32 + i * 4) =
                make_float4(acc[i].x * inv, acc[i].y * inv, acc[i].z * inv, acc[i].w * inv);
    if (sub == 0 && lse_out && r < lse_valid) lse_out[rh] = m + logf(sum);     // [row][head], training only
    __syncthreads();
}

// Phased variant (score items flattened over (row, head, key) -> softmax -> P.V with 8 lanes per
// (row, head)): three CTA barriers, but every thread always has independent items, which wins when the
// context is short (the decode's first ~64 positions) and the K/V rows stream from HBM.
template <int RPT = 4>
__device__ __forceinline__ void attention_phased(const float* __restrict__ sQ, float* __restrict__ sOut,
                                          float* __restrict__ sScore, float* __restrict__ sInv, int lpad,
                                          const float* const* __restrict__ kptr, const float* const* __restrict__ vptr,
                                          const int* __restrict__ nt, int nt_max, float* lse_out = nullptr,
                                          int lse_valid = 0) {
    constexpr int TM = 16 * RPT;
    const int tid = threadIdx.x;
    const float scale = 0.17677669529663687f;   // 1/sqrt(32)
    // phase A: scores
    for (int idx = tid; idx < TM * 2 * nt_max; idx += NT) {
        const int t = idx % nt_max, rh = idx / nt_max, r = rh >> 1, h = rh & 1;
        if (t < nt[r]) {
            const float4* kp = reinterpret_cast<const float4*>(kptr[r] + (long)t * 64 + h * 32);
            const float4* qp = reinterpret_cast<const float4*>(sQ + r * XS + h * 32);
            float4 kk[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) kk[i] = kp[i];
            float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float4 q = qp[i];
                s0 = fmaf(q.x, kk[i].x, s0); s1 = fmaf(q.y, kk[i].y, s1); s2 = fmaf(q.z, kk[i].z, s2); s3 = fmaf(q.w, kk[i].w, s3);
            }
            sScore[rh * lpad + t] = ((s0 + s1) + (s2 + s3)) * scale;
        }
    }
    __syncthreads();
    // phase B: softmax numerators + 1/sum; G = NT / (2 TM) adjacent lanes share one (row, head)
    {
        constexpr int G = NT / (TM * 2);
        const int rh = tid / G, sub = tid % G;
        const int n = nt[rh >> 1];
        float* sc = sScore + rh * lpad;
        float m = -INFINITY;
        for (int t = sub; t < n; t += G) m = fmaxf(m, sc[t]);
#pragma unroll
        for (int o = 1; o < G; o <<= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        float sum = 0.f;
        for (int t = sub; t < n; t += G) {
            const float e = expf(sc[t] - m);
            sc[t] = e;
            sum += e;
        }
#pragma unroll
        for (int o = 1; o < G; o <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        if (sub == 0) {
            sInv[rh] = n > 0 ? 1.0f / sum : 0.f;
            if (lse_out && (rh >> 1) < lse_valid) lse_out[rh] = m + logf(sum);
        }
    }
    __syncthreads();
    // phase C: P.V  (8 threads per (row, head), 4 dims each)
    for (int idx = tid; idx < TM * 2 * 8; idx += NT) {
        const int dq = idx & 7, rh = idx >> 3, r = rh >> 1, h = rh & 1;
        const int n = nt[r];
        const float* sc = sScore + rh * lpad;
        const float* vp = vptr[r] + h * 32 + dq * 4;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        int t = 0;
        for (; t + 4 <= n; t += 4) {
            float4 v[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) v[i] = *reinterpret_cast<const float4*>(vp + (long)(t + i) * 64);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float p = sc[t + i];
                acc.x = fmaf(p, v[i].x, acc.x); acc.y = fmaf(p, v[i].y, acc.y);
                acc.z = fmaf(p, v[i].z, acc.z); acc.w = fmaf(p, v[i].w, acc.w);
            }
        }
        for (; t < n; ++t) {
            const float4 v = *reinterpret_cast<const float4*>(vp + (long)t * 64);
            const float p = sc[t];
            acc.x = fmaf(p, v.x, acc.x); acc.y = fmaf(p, v.y, acc.y); acc.z = fmaf(p, v.z, acc.z); acc.w = fmaf(p, v.w, acc.w);
        }
        const float inv = sInv[rh];
        *reinterpret_cast<float4*>(sOut + r * XS + h * 32 + dq * 4) =
            make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
    }
    __syncthreads();
}


// Absorbed attention (decode).  The cache row t of query row r is the 64-float INPUT x_t of this
// attention (xptr[r] + t*64, written earlier by this CTA); sQ0 / sQ1 hold the folded queries q'_h
// (scale and Wk_h^T already applied), so score_h(t) = q'_h . x_t and ctx_h = sum_t p_h(t) x_t; the
// folded output maps turn (ctx_0, ctx_1) into the attention block's output.  One cache row serves
// both heads.  Outputs overwrite sQ0 / sQ1 with ctx_0 / ctx_1.
template <int RPT = 4>
__device__ __forceinline__ void attention_absorbed(float* __restrict__ sQ0, float* __restrict__ sQ1,
                                                   float* __restrict__ sScore, float* __restrict__ sInv, int lpad,
                                                   const float* const* __restrict__ xptr, const int* __restrict__ nt, int n) {
    constexpr int TM = 16 * RPT;
    const int tid = threadIdx.x;
    // phase A: both heads' scores for item (row, key)
    for (int idx = tid; idx < TM * n; idx += NT) {
        const int r = idx / n, t = idx - r * n;
        if (t < nt[r]) {
            const float4* xp = reinterpret_cast<const float4*>(xptr[r] + (long)t * 64);
            float4 x[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] = xp[i];
            const float4* q0 = reinterpret_cast<const float4*>(sQ0 + r * XS);
            const float4* q1 = reinterpret_cast<const float4*>(sQ1 + r * XS);
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f, b0 = 0.f, b1 = 0.f, b2 = 0.f, b3 = 0.f;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float4 qa = q0[i], qb = q1[i];
                a0 = fmaf(qa.x, x[i].x, a0); a1 = fmaf(qa.y, x[i].y, a1); a2 = fmaf(qa.z, x[i].z, a2); a3 = fmaf(qa.w, x[i].w, a3);
                b0 = fmaf(qb.x, x[i].x, b0); b1 = fmaf(qb.y, x[i].y, b1); b2 = fmaf(qb.z, x[i].z, b2); b3 = fmaf(qb.w, x[i].w, b3);
            }
            sScore[(r * 2) * lpad + t] = (a0 + a1) + (a2 + a3);
            sScore[(r * 2 + 1) * lpad + t] = (b0 + b1) + (b2 + b3);
        }
    }
    __syncthreads();
    // phase B: softmax numerators + 1/sum, G adjacent lanes per (row, head)
    {
        constexpr int G = NT / (TM * 2);
        const int rh = tid / G, sub = tid % G;
        const int m_n = nt[rh >> 1];
        float* sc = sScore + rh * lpad;
        float m = -INFINITY;
        for (int t = sub; t < m_n; t += G) m = fmaxf(m, sc[t]);
#pragma unroll
        for (int o = 1; o < G; o <<= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        float sum = 0.f;
        for (int t = sub; t < m_n; t += G) {
            const float e = expf(sc[t] - m);
            sc[t] = e;
            sum += e;
        }
#pragma unroll
        for (int o = 1; o < G; o <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        if (sub == 0) sInv[rh] = m_n > 0 ? 1.0f / sum : 0.f;
    }
    __syncthreads();
    // phase C: ctx_0, ctx_1 for item (row, 4-column chunk)
    for (int idx = tid; idx < TM * 16; idx += NT) {
        const int r = idx >> 4, dq = idx & 15;
        const int m_n = nt[r];
        const float* s0 = sScore + (r * 2) * lpad;
        const float* s1 = s0 + lpad;
        const float* xp = xptr[r] + dq * 4;
        float4 c0 = make_float4(0.f, 0.f, 0.f, 0.f), c1 = c0;
        int t = 0;
        for (; t + 4 <= m_n; t += 4) {
            float4 x[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) x[i] = *reinterpret_cast<const float4*>(xp + (long)(t + i) * 64);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float p0 = s0[t + i], p1 = s1[t + i];
                c0.x = fmaf(p0, x[i].x, c0.x); c0.y = fmaf(p0, x[i].y, c0.y); c0.z = fmaf(p0, x[i].z, c0.z); c0.w = fmaf(p0, x[i].w, c0.w);
                c1.x = fmaf(p1, x[i].x, c1.x); c1.y = fmaf(p1, x[i].y, c1.y); c1.z = fmaf(p1, x[i].z, c1.z); c1.w = fmaf(p1, x[i].w, c1.w);
            }
        }
        for (; t < m_n; ++t) {
            const float4 x = *reinterpret_cast<const float4*>(xp + (long)t * 64);
            const float p0 = s0[t], p1 = s1[t];
            c0.x = fmaf(p0, x.x, c0.x); c0.y = fmaf(p0, x.y, c0.y); c0.z = fmaf(p0, x.z, c0.z); c0.w = fmaf(p0, x.w, c0.w);
            c1.x = fmaf(p1, x.x, c1.x); c1.y = fmaf(p1, x.y, c1.y); c1.z = fmaf(p1, x.z, c1.z); c1.w = fmaf(p1, x.w, c1.w);
        }
        const float i0 = sInv[r * 2], i1 = sInv[r * 2 + 1];
        *reinterpret_cast<float4*>(sQ0 + r * XS + dq * 4) = make_float4(c0.x * i0, c0.y * i0, c0.z * i0, c0.w * i0);
        *reinterpret_cast<float4*>(sQ1 + r * XS + dq * 4) = make_float4(c1.x * i1, c1.y * i1, c1.z * i1, c1.w * i1);
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
}  // namespace tile
