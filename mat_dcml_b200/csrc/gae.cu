// GAE / returns for sm_100a: one thread per 4 consecutive (env, agent) columns, reverse scan over
// time with 128-bit loads and stores.  HBM-bound: 20 bytes per (t, env, agent) element
// (reward, value, mask read; advantage, return written).
//
// Restates SharedReplayBuffer.compute_returns (mat_src/mat/utils/shared_buffer.py:207-238, GAE
// branch) with ValueNorm.denormalize fused (mat_src/mat/utils/valuenorm.py:32-36,69-79).  Every
// fp32 operation is rounded separately, in numpy's order (-fmad=false), so results are bit-exact.
#include "../../include/matas_b200.h"
#include "common.cuh"

namespace {

struct F4 { float v[4]; };

__device__ __forceinline__ F4 ld4(const float* p, long i, long n) {
    F4 r;
    if (i + 3 < n && ((reinterpret_cast<uintptr_t>(p + i) & 15) == 0)) {
        const float4 x = __ldcs(reinterpret_cast<const float4*>(p + i));
        r.v[0] = x.x; r.v[1] = x.y; r.v[2] = x.z; r.v[3] = x.w;
    } else {
        for (int k = 0; k < 4; ++k) r.v[k] = i + k < n ? p[i + k] : 0.f;
    }
    return r;
}
__device__ __forceinline__ void st4(float* p, long i, long n, const F4& r) {
    if (i + 3 < n && ((reinterpret_cast<uintptr_t>(p + i) & 15) == 0)) {
        __stcs(reinterpret_cast<float4*>(p + i), make_float4(r.v[0], r.v[1], r.v[2], r.v[3]));
    } else {
        for (int k = 0; k < 4; ++k) if (i + k < n) p[i + k] = r.v[k];
    }
}

__global__ void gae_kernel(const float* __restrict__ rewards, const float* __restrict__ values,
                           const float* __restrict__ masks, const float* __restrict__ vn, float gamma, float gl, int T,
                           long n, float* __restrict__ adv, float* __restrict__ ret) {
    const long i = ((long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= n) return;
    float mean = 0.f, sd = 1.f;
    const bool den = vn != nullptr;
    if (den) {                                     // valuenorm.py:32-36
        const float d = fmaxf(vn[2], 1e-5f);
        mean = vn[0] / d;
        const float msq = vn[1] / d;
        sd = sqrtf(fmaxf(msq - mean * mean, 1e-2f));
    }
    float g[4] = {0.f, 0.f, 0.f, 0.f};
    F4 vnext = ld4(values + (long)T * n, i, n);
    if (den) for (int k = 0; k < 4; ++k) vnext.v[k] = vnext.v[k] * sd + mean;
    for (int t = T - 1; t >= 0; --t) {
        const F4 r = ld4(rewards + (long)t * n, i, n);
        const F4 m = ld4(masks + (long)(t + 1) * n, i, n);
        F4 v = ld4(values + (long)t * n, i, n);
        F4 a, rt;
        for (int k = 0; k < 4; ++k) {
            const float vd = den ? v.v[k] * sd + mean : v.v[k];
            const float delta = (r.v[k] + (gamma * vnext.v[k]) * m.v[k]) - vd;     // shared_buffer.py:217-219
            g[k] = delta + ((gl * m.v[k]) * g[k]);                                  // :220
            a.v[k] = g[k];
            rt.v[k] = g[k] + vd;                                                    // :227
            vnext.v[k] = vd;
        }
        st4(adv + (long)t * n, i, n, a);
        st4(ret + (long)t * n, i, n, rt);
    }
}

}  // namespace

extern "C" int matas_gae(const float* rewards, const float* value_preds, const float* masks, const float* vn_stats,
                         double gamma, double gae_lambda, int T, long n, float* advantages, float* returns, void* stream) {
    ARG_CHECK(rewards && value_preds && masks && advantages && returns, "null argument");
    ARG_CHECK(T > 0 && n > 0, "bad shape T=%d n=%ld", T, n);
    const long threads = (n + 3) / 4;
    const unsigned grid = (unsigned)((threads + 127) / 128);
    // numpy multiplies fp32 arrays by python floats: gamma and gamma*lambda are rounded to fp32 once
    gae_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(rewards, value_preds, masks, vn_stats, (float)gamma,
                                                       (float)(gamma * gae_lambda), T, n, advantages, returns);
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}
