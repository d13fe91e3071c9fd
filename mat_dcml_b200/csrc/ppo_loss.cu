// PPO surrogate + clipped value loss of MATTrainer.ppo_update / cal_value_loss, and their gradients with respect to the
// new log-probs and values, in ONE launch (sm_100a).
//
// The reference writes the losses as ~30 element-wise torch expressions over (M * n_agent, 1) vectors
// (mat_src/mat/algorithms/mat/mat_trainer.py:54-94 cal_value_loss, :121-137 surrogate; huber_loss / mse_loss of
// mat_src/mat/utils/util.py:23-29) and lets autograd walk them back: ~120 tiny launches per minibatch between the model's
// forward and backward, during which the GPU waits for the host to issue them.  Here each element's loss terms and the
// closed-form derivatives autograd would produce are computed together:
//     r = exp(lp - lp_old), s1 = r A, s2 = clamp(r, 1 - c, 1 + c) A, m = min(s1, s2)
//     value target t = (R - mean) / std (ValueNorm.normalize, valuenorm.py:58-67) or R
//     e_o = t - v, e_c = t - (v_old + clamp(v - v_old, -c, c)), l = max(H(e_o), H(e_c)) (or H(e_o)), H = huber or e^2 / 2
// torch.min / torch.max send half of the gradient to each argument on an exact tie and clamp passes it on the closed interval;
// huber_loss's second mask is (e > d), not (|e| > d), as in the reference: errors below -d contribute nothing.
// Outputs: d loss / d lp and d loss / d v (already scaled by mask / denominator and by the value-loss coefficient), the
// importance weights, and the two masked loss SUMS (float64 atomics), which the caller divides by its denominators.
#include "../../include/matas_b200.h"
#include "common.cuh"

namespace {

constexpr int F_HUBER = 1, F_CLIPPED = 2, F_POLICY_MASK = 4, F_VALUE_MASK = 8;

__device__ __forceinline__ void loss_term(float e, float delta, bool huber, float& l, float& dl) {   // H(e), H'(e)
    if (huber) {
        const float a = fabsf(e) <= delta ? 1.f : 0.f, b = e > delta ? 1.f : 0.f;
        l = a * (e * e) / 2.f + b * delta * (fabsf(e) - delta / 2.f);
        dl = a * e + b * delta;
    } else {
        l = e * e / 2.f;
        dl = e;
    }
}

__global__ void __launch_bounds__(256)
ppo_loss_kernel(long n, const float* __restrict__ lp, const float* __restrict__ lp_old, const float* __restrict__ adv,
                const float* __restrict__ active, const float* __restrict__ values, const float* __restrict__ v_old,
                const float* __restrict__ ret, const float* __restrict__ vn, const float* __restrict__ den, float lo, float hi,
                float clip, float delta, float value_coef, int flags, float* __restrict__ g_lp, float* __restrict__ g_v,
                float* __restrict__ ratio, double* __restrict__ sums) {
    const float inv_p = 1.0f / den[0], inv_v = 1.0f / den[1];
    float mean = 0.f, sd = 1.f;
    if (vn) { mean = vn[0]; sd = vn[1]; }
    float pol = 0.f, val = 0.f;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const float w = active ? active[i] : 1.f;
        // ---- surrogate --------------------------------------------------------------------------------------------
        const float A = adv[i];
        const float r = expf(lp[i] - lp_old[i]);
        ratio[i] = r;
        const float s1 = r * A, s2 = fminf(fmaxf(r, lo), hi) * A;
        const float d1 = A * r, d2 = (r >= lo && r <= hi) ? A * r : 0.f;
        const float dm = s1 < s2 ? d1 : (s2 < s1 ? d2 : 0.5f * (d1 + d2));
        const float wp = (flags & F_POLICY_MASK) ? w : 1.f;
        pol += fminf(s1, s2) * wp;
        g_lp[i] = -(wp * inv_p) * dm;
        // ---- value loss -------------------------------------------------------------------------------------------
        const float v = values[i], vo = v_old[i];
        const float diff = v - vo;
        const float v_clip = vo + fminf(fmaxf(diff, -clip), clip);
        const float t = vn ? (ret[i] - mean) / sd : ret[i];
        float l_o, d_o, l_c, d_c;
        loss_term(t - v, delta, flags & F_HUBER, l_o, d_o);
        loss_term(t - v_clip, delta, flags & F_HUBER, l_c, d_c);
        d_o = -d_o;                                                   // d e_o / d v = -1
        d_c = (diff >= -clip && diff <= clip) ? -d_c : 0.f;           // d e_c / d v = -1 inside the clip interval
        float l = l_o, dv = d_o;
        if (flags & F_CLIPPED) {
            l = fmaxf(l_o, l_c);
            dv = l_o > l_c ? d_o : (l_c > l_o ? d_c : 0.5f * (d_o + d_c));
        }
        const float wv = (flags & F_VALUE_MASK) ? w : 1.f;
        val += l * wv;
        g_v[i] = value_coef * (wv * inv_v) * dv;
    }
    __shared__ float s_red[2][8];
    pol = warp_sum(pol); val = warp_sum(val);
    if ((threadIdx.x & 31) == 0) { s_red[0][threadIdx.x >> 5] = pol; s_red[1][threadIdx.x >> 5] = val; }
    __syncthreads();
    if (threadIdx.x < 2) {
        double s = 0.0;
        for (int k = 0; k < 8; ++k) s += (double)s_red[threadIdx.x][k];
        atomicAdd(sums + threadIdx.x, s);
    }
}

}  // namespace

extern "C" int matas_ppo_loss(long n, const float* logp, const float* old_logp, const float* adv, const float* active,
                              const float* values, const float* value_preds, const float* returns, const float* vn_mean_std,
                              const float* denominators, float clip_param, float huber_delta, float value_loss_coef, int use_huber,
                              int use_clipped_value_loss, int policy_active_masks, int value_active_masks, float* g_logp,
                              float* g_values, float* ratio, double* sums, void* stream) {
    ARG_CHECK(n > 0 && logp && old_logp && adv && values && value_preds && returns && denominators && g_logp && g_values && ratio && sums,
              "null pointer or n = %ld", n);
    ARG_CHECK(active || !(policy_active_masks || value_active_masks), "active masks requested without an active array");
    const int flags = (use_huber ? F_HUBER : 0) | (use_clipped_value_loss ? F_CLIPPED : 0) | (policy_active_masks ? F_POLICY_MASK : 0) |
                      (value_active_masks ? F_VALUE_MASK : 0);
    // the clamp bounds as torch forms them: Python doubles 1 -/+ clip, rounded to the tensor's float32
    const float lo = (float)(1.0 - (double)clip_param), hi = (float)(1.0 + (double)clip_param);
    const unsigned grid = (unsigned)((n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184);      // at most 8 CTAs per SM
    CUDA_TRY(cudaMemsetAsync(sums, 0, 2 * sizeof(double), (cudaStream_t)stream));
    ppo_loss_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(n, logp, old_logp, adv, active, values, value_preds, returns, vn_mean_std,
                                                            denominators, lo, hi, clip_param, huber_delta, value_loss_coef, flags, g_logp,
                                                            g_values, ratio, sums);
    CUDA_TRY(cudaGetLastError());
    return MATAS_OK;
}
