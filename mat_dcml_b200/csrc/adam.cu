// Fused gradient clipping + Adam step on the flat parameter vector for sm_100a.
//
// Replaces, for the one flat fp32 buffer that holds every parameter of the policy,
//     nn.utils.clip_grad_norm_(policy.transformer.parameters(), max_grad_norm)   (mat_src/mat/algorithms/mat/mat_trainer.py:149-152)
//     policy.optimizer.step()  with  torch.optim.Adam(lr, eps, weight_decay)      (mat_trainer.py:154, transformer_policy.py:103-105)
// i.e. ~25 multi-tensor launches over ~110 parameter views, by two launches: a one-CTA norm of the
// 0.6 MB gradient (it also stores the clip coefficient on the device, so nothing returns to the host)
// and one elementwise pass that scales the gradient in place (as clip_grad_norm_ does), applies L2
// weight decay, updates both moments and the parameters.  HBM-trivial: 28 bytes per parameter.
#include "../../include/matas_b200.h"
#include "common.cuh"

namespace {

constexpr int NORM_THREADS = 1024;

// total = ||g||_2 ; coef = min(1, max_norm / (total + 1e-6))   (torch/nn/utils/clip_grad.py semantics)
__global__ void __launch_bounds__(NORM_THREADS, 1)
adam_norm_kernel(const float* __restrict__ g, long n, float max_norm, float* __restrict__ out) {
    __shared__ double red[NORM_THREADS / 32];
    double acc = 0.0;
    const long n4 = n >> 2;
    const float4* g4 = reinterpret_cast<const float4*>(g);
    for (long i = threadIdx.x; i < n4; i += NORM_THREADS) {
        const float4 x = g4[i];
        acc += (double)(x.x * x.x + x.y * x.y) + (double)(x.z * x.z + x.w * x.w);
    }
    for (long i = (n4 << 2) + threadIdx.x; i < n; i += NORM_THREADS) acc += (double)(g[i] * g[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        double v = threadIdx.x < NORM_THREADS / 32 ? red[threadIdx.x] : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (threadIdx.x == 0) {
            const float total = (float)sqrt(v);
            float coef = 1.0f;
            if (max_norm > 0.f) coef = fminf(max_norm / (total + 1e-6f), 1.0f);
            out[0] = total;
            out[1] = coef;
        }
    }
}

struct AdamArgs {
    float beta1, beta2, one_m_beta1, one_m_beta2, eps, weight_decay, step_size, bc2_sqrt;
};

__device__ __forceinline__ void adam_one(float& p, float& g, float& m, float& v, float coef, const AdamArgs& a) {
    g *= coef;                                            // clip_grad_norm_ scales .grad in place
    float gg = g;
    if (a.weight_decay != 0.f) gg = fmaf(a.weight_decay, p, gg);
    m = fmaf(a.one_m_beta1, gg - m, m);                   // exp_avg.lerp_(grad, 1 - beta1)
    v = fmaf(a.one_m_beta2 * gg, gg, v * a.beta2);        // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value = 1 - beta2)
    const float denom = sqrtf(v) / a.bc2_sqrt + a.eps;
    p = fmaf(-a.step_size, m / denom, p);                 // param.addcdiv_(exp_avg, denom, value = -lr / bias_correction1)
}

__global__ void __launch_bounds__(256)
adam_step_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, long n,
                 const float* __restrict__ norm, AdamArgs a) {
    const float coef = norm[1];
    const long i = ((long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= n) return;
    if (i + 3 < n) {
        float4 pp = *reinterpret_cast<float4*>(p + i), gg = *reinterpret_cast<float4*>(g + i);
        float4 mm = *reinterpret_cast<float4*>(m + i), vv = *reinterpret_cast<float4*>(v + i);
        adam_one(pp.x, gg.x, mm.x, vv.x, coef, a);
        adam_one(pp.y, gg.y, mm.y, vv.y, coef, a);
        adam_one(pp.z, gg.z, mm.z, vv.z, coef, a);
        adam_one(pp.w, gg.w, mm.w, vv.w, coef, a);
        *reinterpret_cast<float4*>(p + i) = pp; *reinterpret_cast<float4*>(g + i) = gg;
        *reinterpret_cast<float4*>(m + i) = mm; *reinterpret_cast<float4*>(v + i) = vv;
    } else {
        for (long k = i; k < n; ++k) adam_one(p[k], g[k], m[k], v[k], coef, a);
    }
}

}  // namespace

extern "C" int matas_adam_clip_step(float* flat_params, float* flat_grad, float* exp_avg, float* exp_avg_sq, long n,
                                    double lr, double beta1, double beta2, double eps, double weight_decay,
                                    double max_grad_norm, long step, float* norm_out, void* stream) {
    ARG_CHECK(flat_params && flat_grad && exp_avg && exp_avg_sq && norm_out, "null argument");
    ARG_CHECK(n > 0 && step >= 1, "bad size / step: n=%ld step=%ld", n, step);
    ARG_CHECK(((reinterpret_cast<uintptr_t>(flat_params) | reinterpret_cast<uintptr_t>(flat_grad) |
                reinterpret_cast<uintptr_t>(exp_avg) | reinterpret_cast<uintptr_t>(exp_avg_sq)) & 15) == 0,
              "flat vectors must be 16-byte aligned");
    ARG_CHECK(beta1 >= 0.0 && beta1 < 1.0 && beta2 >= 0.0 && beta2 < 1.0, "betas out of range");
    cudaStream_t st = (cudaStream_t)stream;
    adam_norm_kernel<<<1, NORM_THREADS, 0, st>>>(flat_grad, n, (float)max_grad_norm, norm_out);
    CUDA_TRY(cudaGetLastError());
    // bias corrections as torch.optim.Adam computes them on the host (python floats), rounded to fp32 once
    const double bc1 = 1.0 - pow(beta1, (double)step), bc2 = 1.0 - pow(beta2, (double)step);
    AdamArgs a;
    a.beta1 = (float)beta1; a.beta2 = (float)beta2;
    a.one_m_beta1 = (float)(1.0 - beta1); a.one_m_beta2 = (float)(1.0 - beta2);
    a.eps = (float)eps; a.weight_decay = (float)weight_decay;
    a.step_size = (float)(lr / bc1);
    a.bc2_sqrt = (float)sqrt(bc2);
    const long threads = (n + 3) / 4;
    adam_step_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(flat_params, flat_grad, exp_avg, exp_avg_sq, n,
                                                                       norm_out, a);
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}
