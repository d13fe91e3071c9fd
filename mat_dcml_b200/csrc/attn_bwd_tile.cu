// Register-tiled attention BACKWARD over whole envs (n_agent <= 64), sm_100a: the counterpart of attn_tile.cu.
//
// The reference differentiates SelfAttention (ma_transformer.py:24-67) with torch autograd; here, per env and head,
//   phase 1: each thread owns a 4 x 4 block of (query i, key j) pairs:  s = q_i.k_j and dp = dY_i.v_j over the
//            head's 32 dims (16 LDS.128 + 128 FFMA per 4 dims), then p = exp2(c s - lse_i), ds = p (dp - D_i)
//            -> shared P[i][j], dS[i][j]
//   phase 2: each thread owns 4 rows x 4 dims of one of
//            dQ_i = scale sum_j dS[i][j] k_j     dK_j = scale sum_i dS[i][j] q_i     dV_j = sum_i P[i][j] dY_i
// with the saved log-sum-exp and D = rowsum(dY * Y) of the forward / projection backward.  Several envs share a
// CTA when they are short, so that the 4 x 4 tasks fill the threads.  dQ / dK / dV go to token-major global
// buffers: the projection backward (policy_bwd.cu) and the weight-gradient kernel (bwd_dw_tc.cu) read them there.
#include <stdlib.h>

#include "policy_bwd.h"
#include "tile.cuh"

namespace {

constexpr int BNT = 128;     // maximum threads per CTA (the launch may use fewer: loops stride by blockDim.x)
constexpr int RS = 68;       // row stride of the Q / K / V / dY tiles (floats)
constexpr float SCALE = 0.17677669529663687f;
constexpr float LOG2E = 1.4426950408889634f;

struct BGeom { int Lp, nb, SS; size_t env_floats; };
__host__ __device__ inline BGeom bgeom(int L) {
    BGeom g;
    g.Lp = (L + 3) & ~3;
    g.nb = g.Lp >> 2;
    g.SS = (g.nb & 1) ? g.Lp : g.Lp + 4;                      // (SS / 4) odd: scalar column walks spread over the banks
    g.env_floats = 4 * (size_t)g.Lp * RS + 2 * (size_t)g.Lp * g.SS + 4 * (size_t)g.Lp;
    return g;
}

template <bool CAUSAL>
__global__ void __launch_bounds__(BNT)
attn_bwd_env_kernel(const float* __restrict__ Q, const float* __restrict__ K, const float* __restrict__ V, const float* __restrict__ gY,
                    const float* __restrict__ lse, const float* __restrict__ Dr, float* __restrict__ dQ, float* __restrict__ dK,
                    float* __restrict__ dV, long n_env, int L, int E) {
    extern __shared__ __align__(16) float sm[];
    const BGeom g = bgeom(L);
    const int Lp = g.Lp, nb = g.nb, SS = g.SS;
    const size_t EF = g.env_floats;
    const int tid = threadIdx.x, nthr = blockDim.x;
    const float c2 = SCALE * LOG2E;
    // per env slot e: sQ | sK | sV | sY [Lp][RS], sP | sdS [Lp][SS], lse2[2][Lp], D[2][Lp]
    auto slot = [&](int e) { return sm + (size_t)e * EF; };
    const size_t oK = (size_t)Lp * RS, oV = 2 * oK, oY = 3 * oK, oP = 4 * oK, oS = oP + (size_t)Lp * SS, oL = oS + (size_t)Lp * SS,
                 oD = oL + 2 * (size_t)Lp;
    for (long env0 = (long)blockIdx.x * E; env0 < n_env; env0 += (long)gridDim.x * E) {
        const int ne = (int)min((long)E, n_env - env0);
        // ---- stage Q, K, V, dY (rows >= L zero), lse (log2 units), D ------------------------------------------
        for (int idx = tid; idx < ne * Lp * 16; idx += nthr) {
            const int e = idx / (Lp * 16), rem = idx - e * Lp * 16, r = rem >> 4, c = rem & 15;
            float4 q = make_float4(0.f, 0.f, 0.f, 0.f), k = q, v = q, y = q;
            if (r < L) {
                const long t = (env0 + e) * L + r;
                q = __ldg(reinterpret_cast<const float4*>(Q + t * 64) + c); k = __ldg(reinterpret_cast<const float4*>(K + t * 64) + c);
                v = __ldg(reinterpret_cast<const float4*>(V + t * 64) + c); y = __ldg(reinterpret_cast<const float4*>(gY + t * 64) + c);
            }
            float* s = slot(e);
            *reinterpret_cast<float4*>(s + r * RS + c * 4) = q;
            *reinterpret_cast<float4*>(s + oK + r * RS + c * 4) = k;
            *reinterpret_cast<float4*>(s + oV + r * RS + c * 4) = v;
            *reinterpret_cast<float4*>(s + oY + r * RS + c * 4) = y;
        }
        for (int idx = tid; idx < ne * 2 * Lp; idx += nthr) {
            const int e = idx / (2 * Lp), rem = idx - e * 2 * Lp, h = rem / Lp, r = rem - h * Lp;
            const long t = (env0 + e) * L + r;
            slot(e)[oL + rem] = r < L ? __ldg(lse + t * 2 + h) * LOG2E : 0.f;
            slot(e)[oD + rem] = r < L ? __ldg(Dr + t * 2 + h) : 0.f;
        }
        __syncthreads();
        for (int h = 0; h < 2; ++h) {
            // ---- phase 1: P and dS blocks (queries {qb + nb i} x keys {kb + nb j}) ---------------------------------
            for (int task = tid; task < ne * nb * nb; task += nthr) {
                const int e = task / (nb * nb), rem = task - e * nb * nb, qb = rem / nb, kb = rem - qb * nb;
                const float* s = slot(e);
                // S = q . k and dP = dY . v as packed fp32x2 products (two partial sums per entry: the x|z and y|w dims), one
                // 4 x 4 block at a time so that the doubled accumulators fit the register budget
                float as[4][4], ap[4][4];
                const float* qp = s + qb * RS + h * 32;
                const float* kp = s + oK + kb * RS + h * 32;
                const float* vp = s + oV + kb * RS + h * 32;
                const float* yp = s + oY + qb * RS + h * 32;
#pragma unroll
                for (int pass = 0; pass < 2; ++pass) {
                    const float* ap_ = pass == 0 ? qp : yp;
                    const float* bp_ = pass == 0 ? kp : vp;
                    float sx[4][4], sy[4][4];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) { sx[i][j] = 0.f; sy[i][j] = 0.f; }
#pragma unroll 2
                    for (int dc = 0; dc < 8; ++dc) {
                        float4 a4[4], b4[4];
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            a4[i] = *reinterpret_cast<const float4*>(ap_ + i * nb * RS + dc * 4);
                            b4[i] = *reinterpret_cast<const float4*>(bp_ + i * nb * RS + dc * 4);
                        }
#pragma unroll
                        for (int i = 0; i < 4; ++i)
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                fma2(sx[i][j], sy[i][j], a4[i].x, a4[i].y, b4[j].x, b4[j].y);
                                fma2(sx[i][j], sy[i][j], a4[i].z, a4[i].w, b4[j].z, b4[j].w);
                            }
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            if (pass == 0) as[i][j] = sx[i][j] + sy[i][j];
                            else ap[i][j] = sx[i][j] + sy[i][j];
                        }
                }
                float* sp = slot(e) + oP;
                float* sd = slot(e) + oS;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int qi = qb + nb * i;
                    const float l2 = s[oL + h * Lp + qi], dd = s[oD + h * Lp + qi];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int kj = kb + nb * j;
                        const bool ok = qi < L && kj < L && (!CAUSAL || kj <= qi);
                        const float p = ok ? exp2f(fmaf(as[i][j], c2, -l2)) : 0.f;
                        sp[qi * SS + kj] = p;
                        sd[qi * SS + kj] = p * (ap[i][j] - dd);
                    }
                }
            }
            __syncthreads();
            // ---- phase 2: dQ / dK / dV blocks of 4 rows x 4 dims -----------------------------------------------------
            for (int task = tid; task < ne * 3 * nb * 8; task += nthr) {
                const int e = task / (3 * nb * 8), rem = task - e * 3 * nb * 8, kind = rem / (nb * 8), r2 = rem - kind * nb * 8;
                const int rb = r2 >> 3, db = r2 & 7;
                const float* s = slot(e);
                float acc[4][4];
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
                if (kind == 0) {                       // dQ rows 4rb..: sum over keys t of dS[row][t] k_t
                    const float* sd = s + oS + (rb * 4) * SS;
                    const float* kp = s + oK + h * 32 + db * 4;
                    const int nt = CAUSAL ? min(L, rb * 4 + 4) : L;
                    for (int t = 0; t < nt; ++t) {
                        const float4 k4 = *reinterpret_cast<const float4*>(kp + t * RS);
                        const float w[4] = {sd[t], sd[SS + t], sd[2 * SS + t], sd[3 * SS + t]};
#pragma unroll
                        for (int i = 0; i < 4; ++i) {                 // packed: scalar weight x two dims per instruction
                            fma2(acc[i][0], acc[i][1], w[i], w[i], k4.x, k4.y);
                            fma2(acc[i][2], acc[i][3], w[i], w[i], k4.z, k4.w);
                        }
                    }
                } else {                               // dK / dV rows (keys) 4rb..: sum over queries t of dS|P[t][row] q_t|dY_t
                    const float* sw = s + (kind == 1 ? oS : oP) + rb * 4;
                    const float* xp = s + (kind == 1 ? 0 : oY) + h * 32 + db * 4;
                    const int t0 = CAUSAL ? rb * 4 : 0;
                    for (int t = t0; t < L; ++t) {
                        const float4 w4 = *reinterpret_cast<const float4*>(sw + t * SS);
                        const float4 x4 = *reinterpret_cast<const float4*>(xp + t * RS);
                        const float w[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            fma2(acc[i][0], acc[i][1], w[i], w[i], x4.x, x4.y);
                            fma2(acc[i][2], acc[i][3], w[i], w[i], x4.z, x4.w);
                        }
                    }
                }
                float* out = kind == 0 ? dQ : (kind == 1 ? dK : dV);
                const float mul = kind == 2 ? 1.0f : SCALE;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int r = rb * 4 + i;
                    if (r < L)
                        *reinterpret_cast<float4*>(out + ((env0 + e) * L + r) * 64 + h * 32 + db * 4) =
                            make_float4(acc[i][0] * mul, acc[i][1] * mul, acc[i][2] * mul, acc[i][3] * mul);
                }
            }
            __syncthreads();
        }
    }
}

}  // namespace

bool attn_tile_supported(int L) { return L <= 64; }

int launch_attn_bwd_tile(const float* Q, const float* K, const float* V, const float* gY, const float* lse, const float* D, float* dQ,
                         float* dK, float* dV, long n_env, int L, int causal, cudaStream_t st) {
    const BGeom g = bgeom(L);
    // measured on B200 (12800 x 21 / 4096 x 51): small CTAs with one env each beat packing several envs into a
    // CTA (more resident CTAs hide the shared-memory latency): 64 threads up to 64 phase-1 tasks, else 128
    const int tasks = g.nb * g.nb;
    const int nthr = tasks > 64 ? BNT : 64;
    int E = tasks >= 32 ? 1 : 32 / tasks;
    if (E > 4) E = 4;
    const size_t bytes = (size_t)E * g.env_floats * sizeof(float);
    const long groups = (n_env + E - 1) / E;
    const unsigned grid = (unsigned)(groups < (1L << 20) ? groups : (1L << 20));
    if (causal) {
        CUDA_TRY(cudaFuncSetAttribute(attn_bwd_env_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
        attn_bwd_env_kernel<true><<<grid, nthr, bytes, st>>>(Q, K, V, gY, lse, D, dQ, dK, dV, n_env, L, E);
    } else {
        CUDA_TRY(cudaFuncSetAttribute(attn_bwd_env_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
        attn_bwd_env_kernel<false><<<grid, nthr, bytes, st>>>(Q, K, V, gY, lse, D, dQ, dK, dV, n_env, L, E);
    }
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}
