// Register-tiled attention over one env's tokens (token-parallel passes, n_agent <= 64), sm_100a.
//
// SelfAttention of mat_src/mat/algorithms/mat/algorithm/ma_transformer.py:24-67 for one env at a time: Q, K, V
// rows [L][64] (two heads of 32) are staged in shared memory, then
//   S phase : each thread owns a 4 x 4 block of (query, key) pairs of one head: 8 x (8 LDS.128 + 64 FFMA)
//   softmax : four lanes per (head, query) row; probabilities are written transposed, PT[key][query]
//   PV phase: each thread owns 4 queries x 4 output dims: per key 2 LDS.128 + 16 FFMA
// i.e. about 270 instructions per (query, key) pair and lane less than the half-warp-per-row formulation
// (tile::attention_rows), which spends most of its issue slots on shuffle reductions and per-key softmax updates.
// The output Y [tok][64] (and, in training mode, the natural-log log-sum-exp [tok][2]) goes to global memory and
// is picked up by the tensor-core block kernel as its first GEMM operand.
#include <stdlib.h>

#include "policy_fwd.h"
#include "tile.cuh"

namespace {

constexpr int ANT = 128;     // maximum threads per CTA (loops stride by blockDim.x)
constexpr int RS = 68;       // row stride of the Q / K / V tiles (floats): 16-byte aligned rows, 4-bank skew

struct AttnGeom { int Lp, nb, SS, PS; size_t qk_floats, floats; };
// vr = true: V waits in registers until the score tile is free and then takes its place (less shared memory: 48.8 KB
// instead of 62.6 KB per env at n_agent 51, four resident CTAs per SM instead of three); vr = false: V has its own tile
// (no extra barrier, fewer live registers -- better whenever shared memory does not limit the resident CTAs).
__host__ __device__ inline AttnGeom attn_geom(int L, bool vr) {
    AttnGeom g;
    g.Lp = (L + 3) & ~3;
    g.nb = g.Lp >> 2;
    g.SS = (g.nb & 1) ? g.Lp : g.Lp + 4;                    // S row stride: (SS / 4) odd -> 8 rows x 4 keys hit 32 banks
    g.PS = g.Lp + ((8 - (g.Lp & 15)) & 15);                  // PT row stride: PS = 8 (mod 16), >= Lp
    const size_t qk = 2 * (size_t)g.Lp * RS, pt = 2 * (size_t)g.Lp * g.PS;      // [Q | K], later PT
    const size_t sc = 2 * (size_t)g.Lp * g.SS, vv = (size_t)g.Lp * RS;          // S and V
    g.qk_floats = qk > pt ? qk : pt;
    g.floats = g.qk_floats + (vr ? (sc > vv ? sc : vv) : sc + vv);
    return g;
}

constexpr int VREG = 64 * 16 / ANT;      // float4 registers per thread that hold the V rows of one env (n_agent <= 64)

template <bool CAUSAL, bool VR>
__global__ void __launch_bounds__(ANT)
attn_env_kernel(const float* __restrict__ Q, const float* __restrict__ K, const float* __restrict__ V, float* __restrict__ Y,
                float* __restrict__ lse, long n_env, int L) {
    extern __shared__ __align__(16) float sm[];
    const AttnGeom g = attn_geom(L, VR);
    const int Lp = g.Lp, nb = g.nb, SS = g.SS, PS = g.PS;
    float* sQ = sm;
    float* sK = sm + (size_t)Lp * RS;
    float* sPT = sm;                                          // aliases sQ / sK once the scores are done
    float* sS = sm + g.qk_floats;
    float* sV = VR ? sS : sS + 2 * (size_t)Lp * SS;           // VR: takes the place of the scores once the probabilities are written
    const int tid = threadIdx.x, nthr = blockDim.x;           // nthr == ANT (launch_attn_tile)
    const float scale = 0.17677669529663687f * 1.4426950408889634f;     // log2(e)/sqrt(32): exp2-domain scores
    for (long env = blockIdx.x; env < n_env; env += gridDim.x) {
        const long tok0 = env * L;
        // ---- stage Q, K, V (rows >= L zero) ----------------------------------------------------------------------
        float4 vreg[VREG];
        if (VR) {
#pragma unroll
            for (int i = 0; i < VREG; ++i) {
                const int idx = tid + i * ANT, r = idx >> 4, c = idx & 15;
                vreg[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r < L) vreg[i] = __ldg(reinterpret_cast<const float4*>(V + (tok0 + r) * 64) + c);
            }
        }
        for (int idx = tid; idx < Lp * 16; idx += nthr) {
            const int r = idx >> 4, c = idx & 15;
            float4 q = make_float4(0.f, 0.f, 0.f, 0.f), k = q, v = q;
            if (r < L) {
                q = __ldg(reinterpret_cast<const float4*>(Q + (tok0 + r) * 64) + c);
                k = __ldg(reinterpret_cast<const float4*>(K + (tok0 + r) * 64) + c);
                if (!VR) v = __ldg(reinterpret_cast<const float4*>(V + (tok0 + r) * 64) + c);
            }
            *reinterpret_cast<float4*>(sQ + r * RS + c * 4) = q;
            *reinterpret_cast<float4*>(sK + r * RS + c * 4) = k;
            if (!VR) *reinterpret_cast<float4*>(sV + r * RS + c * 4) = v;
        }
        __syncthreads();
        // ---- S = scale * Q K^T: thread block = queries {qb + nb*i} x keys {kb + nb*j} of head h ----------------
        for (int task = tid; task < 2 * nb * nb; task += nthr) {
            const int h = task / (nb * nb), rem = task - h * nb * nb, qb = rem / nb, kb = rem - qb * nb;
            float acc[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
            const float* qp = sQ + qb * RS + h * 32;
            const float* kp = sK + kb * RS + h * 32;
#pragma unroll 2
            for (int dc = 0; dc < 8; ++dc) {
                float4 q4[4], k4[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) q4[i] = *reinterpret_cast<const float4*>(qp + i * nb * RS + dc * 4);
#pragma unroll
                for (int j = 0; j < 4; ++j) k4[j] = *reinterpret_cast<const float4*>(kp + j * nb * RS + dc * 4);
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        acc[i][j] = fmaf(q4[i].x, k4[j].x, acc[i][j]); acc[i][j] = fmaf(q4[i].y, k4[j].y, acc[i][j]);
                        acc[i][j] = fmaf(q4[i].z, k4[j].z, acc[i][j]); acc[i][j] = fmaf(q4[i].w, k4[j].w, acc[i][j]);
                    }
            }
            float* sp = sS + (size_t)h * Lp * SS;
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) sp[(qb + nb * i) * SS + kb + nb * j] = acc[i][j] * scale;
        }
        __syncthreads();
        // ---- softmax rows: 4 lanes per (head, query); normalised probabilities -> PT[h][key][query] ----------------
        {
            const int grp = tid >> 2, l4 = tid & 3;
            const unsigned gmask = 0xfu << (tid & 28);
            const int n_it = (2 * L + nthr / 4 - 1) / (nthr / 4);
            for (int it = 0; it < n_it; ++it) {
                const int rt = it * (nthr / 4) + grp;
                const bool on = rt < 2 * L;
                const int h = on ? rt / L : 0, qi = on ? rt - h * L : 0;
                const int nk = CAUSAL ? qi + 1 : L;
                const float* sr = sS + ((size_t)h * Lp + qi) * SS;
                float m = -INFINITY;
                for (int k = l4; k < nk; k += 4) m = fmaxf(m, sr[k]);
                m = fmaxf(m, __shfl_xor_sync(gmask, m, 1));
                m = fmaxf(m, __shfl_xor_sync(gmask, m, 2));
                float sum = 0.f;
                for (int k = l4; k < nk; k += 4) sum += exp2f(sr[k] - m);
                sum += __shfl_xor_sync(gmask, sum, 1);
                sum += __shfl_xor_sync(gmask, sum, 2);
                if (on) {
                    const float inv = 1.0f / sum;
                    float* pc = sPT + (size_t)h * Lp * PS + qi;
                    for (int k = l4; k < Lp; k += 4) pc[k * PS] = k < nk ? exp2f(sr[k] - m) * inv : 0.f;
                    if (lse && l4 == 0) lse[(tok0 + qi) * 2 + h] = m * 0.6931471805599453f + logf(sum);
                }
            }
        }
        __syncthreads();
        if (VR) {      // V: registers -> the tile the scores occupied
#pragma unroll
            for (int i = 0; i < VREG; ++i) {
                const int idx = tid + i * ANT, r = idx >> 4, c = idx & 15;
                if (r < Lp) *reinterpret_cast<float4*>(sV + r * RS + c * 4) = vreg[i];
            }
            __syncthreads();
        }
        // ---- Y = P V: thread block = queries 4qb..4qb+3 x dims 4db..4db+3 of head h -------------------------------
        for (int task = tid; task < 2 * nb * 8; task += nthr) {
            const int h = task / (nb * 8), rem = task - h * nb * 8, qb = rem >> 3, db = rem & 7;
            float acc[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
            const float* pp = sPT + (size_t)h * Lp * PS + qb * 4;
            const float* vp = sV + h * 32 + db * 4;
            const int nt = CAUSAL ? min(L, qb * 4 + 4) : L;
#pragma unroll 4
            for (int t = 0; t < nt; ++t) {
                const float4 p = *reinterpret_cast<const float4*>(pp + t * PS);
                const float4 v = *reinterpret_cast<const float4*>(vp + t * RS);
                const float pv[4] = {p.x, p.y, p.z, p.w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    acc[i][0] = fmaf(pv[i], v.x, acc[i][0]); acc[i][1] = fmaf(pv[i], v.y, acc[i][1]);
                    acc[i][2] = fmaf(pv[i], v.z, acc[i][2]); acc[i][3] = fmaf(pv[i], v.w, acc[i][3]);
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int qi = qb * 4 + i;
                if (qi < L)
                    *reinterpret_cast<float4*>(Y + (tok0 + qi) * 64 + h * 32 + db * 4) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
            }
        }
        __syncthreads();
    }
}

}  // namespace

bool attn_tile_supported(int L) { return L <= 64; }

static bool attn_use_tc() {
    static int use_tc = -1;          // MATAS_ATTN_TC=0 selects the register-tiled CUDA-core kernel below (A/B switch)
    if (use_tc < 0) { const char* e = getenv("MATAS_ATTN_TC"); use_tc = e ? atoi(e) : 1; }
    return use_tc != 0;
}
// forward passes can take the separate attention kernel up to n_agent 256 (tensor-core kernel with key tiles)
bool attn_fwd_supported(int L) { return L <= 64 || (attn_use_tc() && L <= 256); }

int launch_attn_tile(const float* Q, const float* K, const float* V, float* Y, float* lse, long n_env, int L, int causal,
                     cudaStream_t st) {
    // tensor-core kernel (attn_tc.cu) unless MATAS_ATTN_TC=0 selects this register-tiled CUDA-core kernel (A/B switch)
    static int n_sm = 0;
    if (n_sm == 0) {
        int dev = 0;
        CUDA_TRY(cudaGetDevice(&dev));
        CUDA_TRY(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
    }
    if (attn_use_tc()) return launch_attn_tc(Q, K, V, Y, lse, n_env, L, causal, n_sm, st);
    ARG_CHECK(L <= 64, "attn_env_kernel: n_agent %d > 64", L);
    // V in registers only where it buys a resident CTA and shared memory is what limits them
    const size_t smem_sm = 227 * 1024, b0 = attn_geom(L, false).floats * sizeof(float) + 1024,
                 b1 = attn_geom(L, true).floats * sizeof(float) + 1024;
    static int vr_env = -1;                             // tuning override: MATAS_AT_VREG = 0 / 1
    if (vr_env < 0) { const char* e = getenv("MATAS_AT_VREG"); vr_env = e ? atoi(e) : 2; }
    const bool vr = vr_env == 2 ? (smem_sm / b0 < 4 && smem_sm / b1 > smem_sm / b0) : vr_env != 0;
    const size_t bytes = attn_geom(L, vr).floats * sizeof(float);
    const unsigned grid = (unsigned)(n_env < (1L << 30) ? n_env : (1L << 30));
#define ATTN_LAUNCH(C, R)                                                                                              \
    do {                                                                                                               \
        CUDA_TRY(cudaFuncSetAttribute(attn_env_kernel<C, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)); \
        attn_env_kernel<C, R><<<grid, ANT, bytes, st>>>(Q, K, V, Y, lse, n_env, L);                                    \
    } while (0)
    if (causal) { if (vr) ATTN_LAUNCH(true, true); else ATTN_LAUNCH(true, false); }
    else { if (vr) ATTN_LAUNCH(false, true); else ATTN_LAUNCH(false, false); }
#undef ATTN_LAUNCH
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}
