/*
 * matas_b200.h -- C ABI of the B200-native MAT-AS decision-and-rollout hot path.
 *
 * The reference (brithbutter/mat-dcml) is pure Python and has no FFI; its boundary for this
 * path is the duck-typed policy / vec-env / buffer contract cited per function below (paths are
 * relative to the reference root).  A Python host binds these entry points with ctypes
 * (mat_dcml_b200/_cabi.py; INTEGRATION.md shows the stub a reference maintainer would add).
 *
 * Conventions
 *  - every function returns 0 on success, non-zero otherwise; matas_last_error() then returns a
 *    thread-local, human-readable message.  No exception crosses the boundary.
 *  - all array arguments are DEVICE pointers to contiguous fp32 (or the stated type) unless the
 *    name ends in _host; they are borrowed for the duration of the call.  `stream` is a
 *    cudaStream_t passed as void*; work is enqueued, not synchronised.
 *  - handles belong to the CUDA device that was current at creation, are not thread-safe and
 *    must be used with that device current.  Functions that take a handle may allocate device
 *    scratch (cudaMalloc) the first time a larger batch is seen.
 *  - layouts are env-major / agent-minor, exactly the reference's flattened (B*L, dim) arrays.
 */
#ifndef MATAS_B200_H
#define MATAS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MATAS_MODE_SEMI 0     /* 'Semi_Discrete': L-1 binary agents + 1 Gaussian agent (every DCML config) */
#define MATAS_MODE_DISCRETE 1 /* 'Discrete': L binary agents */

const char* matas_last_error(void);
int matas_version(void);

/* ------------------------------------------------------------------------------------------
 * Policy: replaces MultiAgentTransformer (mat_src/mat/algorithms/mat/algorithm/ma_transformer.py:233-339)
 * behind TransformerPolicy (.../transformer_policy.py:9-255).
 * ------------------------------------------------------------------------------------------ */
typedef struct matas_policy matas_policy;

typedef struct {
    int n_agent;   /* L = workers + 1 */
    int obs_dim;   /* 7 for DCML (DCML_ENVs/DCML_utils/DCML_Config.py:13-17); <= 8 supported */
    int state_dim; /* share_obs width W+2; only sizes the unused encoder.state_encoder parameters */
    int act_dim;   /* 2 */
    int n_block;   /* 2  (mat_src/mat/config.py:305) */
    int n_embd;    /* 64 (config.py:306) */
    int n_head;    /* 2  (config.py:307) */
    int mode;      /* MATAS_MODE_* ; transformer_policy.py:27-37 */
} matas_model_cfg;

/* transformer_policy.py:82-86 (model construction). */
int matas_policy_create(const matas_model_cfg* cfg, matas_policy** out);
int matas_policy_destroy(matas_policy* p);

/* Number of trainable parameters, i.e. the length of the flat vector below (151,469 at L=101). */
long matas_policy_param_count(const matas_policy* p);

/* Load parameters from ONE flat fp32 device vector holding every nn.Parameter of the reference
 * model in state_dict order (mask buffers excluded).  Replaces load_state_dict /
 * the implicit sharing of module parameters (transformer_policy.py:246-248); call again after
 * every optimizer step. */
int matas_policy_set_weights(matas_policy* p, const float* flat_params, void* stream);

/* MultiAgentTransformer.get_values, ma_transformer.py:331-339 (encoder only).
 * obs [B*L*obs_dim] -> values [B*L]. */
int matas_policy_get_values(matas_policy* p, const float* obs, long B, float* values, void* stream);

/* MultiAgentTransformer.get_actions, ma_transformer.py:298-329, i.e. encoder +
 * semi_discrete_autoregreesive_act / discrete_autoregreesive_act
 * (mat_src/mat/algorithms/utils/transformer_act.py:30-100, 131-173), including the stride-block
 * deterministic schedule.
 *   ava    [B*L*2] availability (1/0) or NULL
 *   u, z   injected randomness of the stochastic branch: u [B*L] uniforms in [0,1), z [B*2]
 *          standard normals; NULL = draw them in-kernel from Philox(seed, env_offset+env, step)
 *   out    values [B*L], actions [B*L] (0/1 as float; the Gaussian agent's real value),
 *          logp [B*L] */
int matas_policy_get_actions(matas_policy* p, const float* obs, const float* ava, long B, int deterministic,
                             int stride, uint64_t seed, uint32_t step, long env_offset, const float* u,
                             const float* z, float* values, float* actions, float* logp, void* stream);

/* MultiAgentTransformer.forward, ma_transformer.py:257-295, i.e. encoder +
 * semi_discrete_parallel_act / discrete_parallel_act (transformer_act.py:103-129, 176-189).
 * actions [M*L] -> values [M*L], logp [M*L], entropy [M*L] (per agent; the masked mean of
 * transformer_policy.py:212-215 is the caller's). */
int matas_policy_evaluate(matas_policy* p, const float* obs, const float* ava, const float* actions, long M,
                          float* values, float* logp, float* entropy, void* stream);

/* MultiAgentTransformer.get_actions (ma_transformer.py:298-329) in two parts, for callers whose observations arrive from
 * the host in slices (ma_transformer.py:304-307 copies them in one piece): matas_policy_encode_slice runs the encoder
 * for the envs [env0, env0 + n_env) of a batch of B_total envs (obs / values point at the slice's first row; the first
 * slice must be env0 = 0) and matas_policy_decode, once every slice is in, the autoregressive act of the whole batch
 * (arguments as matas_policy_get_actions).  Together they compute exactly what matas_policy_get_actions computes. */
int matas_policy_encode_slice(matas_policy* p, const float* obs, long env0, long n_env, long B_total, float* values,
                              void* stream);
int matas_policy_decode(matas_policy* p, const float* ava, long B, int deterministic, int stride, uint64_t seed,
                        uint32_t step, long env_offset, const float* u, const float* z, float* actions, float* logp,
                        void* stream);

/* Training-mode forward + backward of MultiAgentTransformer.forward, replacing torch autograd in
 * MATTrainer.ppo_update (mat_src/mat/algorithms/mat/mat_trainer.py:120-147).
 * matas_policy_evaluate_train = matas_policy_evaluate that also keeps the activations;
 * matas_policy_evaluate_backward must directly follow it with the same batch: given dLoss/dvalues,
 * dLoss/dlogp, dLoss/dentropy [M*L] it ACCUMULATES dLoss/dparams into flat_grad (fp32, laid out like
 * the flat parameter vector; unused parameters keep zero).  n_agent <= 256. */
int matas_policy_evaluate_train(matas_policy* p, const float* obs, const float* ava, const float* actions, long M,
                                float* values, float* logp, float* entropy, void* stream);
int matas_policy_evaluate_backward(matas_policy* p, const float* obs, const float* ava, const float* actions, long M,
                                   const float* d_values, const float* d_logp, const float* d_entropy,
                                   float* flat_grad, void* stream);

/* Encoder.forward, ma_transformer.py:141-154: obs -> values [B*L] and obs_rep [B*L*64]. */
int matas_policy_encode(matas_policy* p, const float* obs, long B, float* values, float* obs_rep, void* stream);

/* The reference's second-level entry points, which start from obs_rep:
 * semi_discrete_autoregreesive_act / discrete_autoregreesive_act (transformer_act.py:30-100, 131-173);
 * arguments as matas_policy_get_actions. */
int matas_policy_act_autoregressive(matas_policy* p, const float* obs_rep, const float* ava, long B, int deterministic,
                                    int stride, uint64_t seed, uint32_t step, long env_offset, const float* u,
                                    const float* z, float* actions, float* logp, void* stream);

/* semi_discrete_parallel_act / discrete_parallel_act (transformer_act.py:103-129, 176-189). */
int matas_policy_act_parallel(matas_policy* p, const float* obs_rep, const float* ava, const float* actions, long M,
                              float* logp, float* entropy, void* stream);

/* Allocation-free variant of matas_policy_act_autoregressive (no counterpart in the reference).
 * For callers that must not allocate device memory after start-up -- e.g. a decision server sharing
 * the GPU with another framework's caching allocator, or a latency-critical loop that cannot take
 * the device-wide synchronisation of cudaMalloc/cudaFree.  Its scratch (cross-attention queries +
 * decode K/V cache) is a fixed 5 MiB array inside the library image, resident on every GPU that
 * loads the library until the process exits.  Limit: B * n_agent <= 2048 (env, agent) pairs per
 * call; larger calls fail with an error.  The first call on each device resolves the array's device
 * address (make it before any stream capture).  All calls on a device share the array: calls that
 * overlap on the GPU race, so serialise them on one stream; nothing needs clearing between calls.
 * Arguments and results as matas_policy_act_autoregressive. */
int matas_policy_act_autoregressive_static(matas_policy* p, const float* obs_rep, const float* ava, long B,
                                           int deterministic, int stride, uint64_t seed, uint32_t step,
                                           long env_offset, const float* u, const float* z, float* actions,
                                           float* logp, void* stream);

/* Measurement hooks (no counterpart in the reference): when enabled, every launch of the
 * autoregressive decode kernel is bracketed by CUDA events on its stream; matas_policy_decode_time
 * waits for them and returns the summed device time and the launch count since enabling. */
int matas_policy_timing(matas_policy* p, int enable);
int matas_policy_decode_time(matas_policy* p, double* ms_total, long* launches);
/* Minibatch gather: replaces the per-array fancy indexing of SharedReplayBuffer.feed_forward_generator_transformer
 * (mat_src/mat/utils/shared_buffer.py:265-314).  For each of n_arrays (<= 12) device arrays: dst[a][i] = src[a][idx[i]]
 * for i < n_rows, a row being width[a] floats, written repeat[a] times back to back (share_obs: one (W+2) row per
 * (t, env) sample replicated n_agent times, shared_buffer.py:265-266).  src / dst / width / repeat are HOST arrays of
 * device pointers / ints; idx is a device int64 vector (the slice of torch.randperm, shared_buffer.py:260-261). */
int matas_gather_rows(int n_arrays, const float* const* src, float* const* dst, const int* width, const int* repeat,
                      const long* idx, long n_rows, void* stream);

/* PPO losses and their gradients in one launch: replaces the element-wise expressions of MATTrainer.ppo_update
 * (mat_src/mat/algorithms/mat/mat_trainer.py:121-137: importance weights, clipped surrogate, masked mean) and
 * cal_value_loss (:54-94: clipped value prediction, ValueNorm-normalised target, huber_loss / mse_loss of
 * mat_src/mat/utils/util.py:23-29, max of clipped and unclipped loss, masked mean), and the autograd pass through them.
 * All arrays are device vectors of n = M * n_agent floats; `active` may be NULL when neither mask flag is set;
 * vn_mean_std = {mean, sqrt(var)} of ValueNorm.running_mean_var() after its update (valuenorm.py:32-36,58-67), or NULL
 * (use_valuenorm False); denominators = {policy denominator, value denominator} (active_masks.sum() or the element count:
 * mat_trainer.py:88-91,127-132), device floats so that no host read is needed.  Outputs: g_logp = d loss / d logp and
 * g_values = d loss / d values for loss = policy_loss + value_loss_coef * value_loss (the entropy term is linear in the
 * entropy the policy returns), ratio = the importance weights, sums = {sum of min(surr1, surr2) * mask, sum of
 * value loss * mask} in float64: policy_loss = -sums[0] / denominators[0], value_loss = sums[1] / denominators[1]. */
int matas_ppo_loss(long n, const float* logp, const float* old_logp, const float* adv, const float* active, const float* values,
                   const float* value_preds, const float* returns, const float* vn_mean_std, const float* denominators,
                   float clip_param, float huber_delta, float value_loss_coef, int use_huber, int use_clipped_value_loss,
                   int policy_active_masks, int value_active_masks, float* g_logp, float* g_values, float* ratio, double* sums,
                   void* stream);

/* Philox4x32-10 exactly as the kernels draw it (csrc/common.cuh; no counterpart in the reference, whose Python `random` /
 * `numpy.random` streams cannot be replayed in parallel -- see DESIGN.md "Random streams"): n device rows of
 * {counter[4], key[2]} -> n rows of 4 output words.  Test hook for the Random123 known-answer vectors. */
int matas_philox4x32_10(const uint32_t* counter_key, uint32_t* out, long n, void* stream);

/* Measured fp32 FFMA throughput of the current device in TFLOP/s (no counterpart in the reference;
 * the roofline denominator of the SIMT kernels, which MEASURED_PEAKS.json does not hold). */
int matas_fp32_peak(int iters, double* tflops);

/* ------------------------------------------------------------------------------------------
 * Vectorised DCML simulator: replaces Env.step / Env.reset
 * (DCML_BID_FIRST_MA_ENV_SingleProcess.py:57-144, 157-274), Worker.process / bid
 * (DCML_Worker_TIMESLOT_MultiProcess.py:46-135), Master.reset / get_workload (DCML_Master.py:35-56)
 * and ShareDummyVecEnv (mat_src/mat/envs/env_wrappers.py:713-744) for B envs at once.
 * Random quantities are the named Philox draws documented in DESIGN.md ("Random streams").
 * ------------------------------------------------------------------------------------------ */
typedef struct dcml_env dcml_env;

typedef struct {
    long n_envs;
    int n_workers;     /* W; agents L = W+1 */
    uint64_t seed;
    long env_offset;   /* global index of env 0 (rank * n_envs when sharded) */
    int double_reset;  /* 1 = ShareDummyVecEnv's second reset on done (env_wrappers.py:730-736) */
    int shannon;       /* 1 = Shannon link rates (reset(shannon_enable=True): DCML_ENVs/DCML_Basic_Env.py:18-27,
                          DCML_Master.py:41-45, Shannon.py:14-21); share_obs rows are then 2W+2 wide
                          [Rn, Cn, upload/1e7 (W), download/1e7 (W)], …SingleProcess.py:248-250 */
    int fixed;         /* 1 = Env(fixed=True), …SingleProcess.py:58-62: the actions are ignored, every available worker is
                          selected and K = floor(0.7 N) */
    int standalone;    /* 1 = every step takes the step(standalone=True) branch, …SingleProcess.py:81-92: the task runs
                          uncoded on worker 0 (K = N = 1) and the reward is scaled by 1.5 */
} dcml_env_cfg;

/* Env.__init__ (…SingleProcess.py:16-55): base_workload_host is the workloads.txt table,
 * [W][20] float64 on the HOST. */
int dcml_env_create(const dcml_env_cfg* cfg, const double* base_workload_host, dcml_env** out);
int dcml_env_destroy(dcml_env* e);

/* Env.reset: obs [B*L*7], share_obs [B*(W+2)] ([B*(2W+2)] in Shannon mode; ONE row per env; the reference replicates it over
 * the L agents, …SingleProcess.py:270), ava [B*L*2]. */
int dcml_env_reset(dcml_env* e, float* obs, float* share_obs, float* ava, void* stream);

/* Env.reset(binary=True) (…SingleProcess.py:164-172): the task size enters share_obs as the 32-bit strings of R and C
 * (most significant bit first) instead of the two normalised values, so a share_obs row is 62 floats wider
 * ([B*(W+64)], Shannon mode [B*(2W+64)]), and observation features 0 / 1 of every agent are that row's first two entries
 * (the two top bits of R, :203).  Only this call is binary: the reset inside a step is not (:139), as in the reference. */
int dcml_env_reset_binary(dcml_env* e, float* obs, float* share_obs_bits, float* ava, void* stream);

/* Env.step: actions [B*L] -> next obs/share_obs/ava plus reward [B], done [B] (uint8),
 * delay [B], payment [B] (the info dict, …SingleProcess.py:143); the reference broadcasts
 * reward/done over the L agents (DCML_ENVs/DCML_Basic_Env.py:9-13). */
int dcml_env_step(dcml_env* e, const float* actions, float* obs, float* share_obs, float* ava, float* reward,
                  uint8_t* done, float* delay, float* payment, void* stream);
/* The same step for the envs [env0, env0 + n_env) only (all pointers are those of the FULL arrays): a host-facing wrapper
 * issues the batch as a few slices so that the device-to-host copy of one slice's outputs runs under the next slice's
 * kernel (env_wrappers.py:722-739 returns host arrays).  A step is complete when every env was covered exactly once;
 * `first` = 1 on the slice that opens a step (it advances the step counter of the named random draws). */
int dcml_env_step_slice(dcml_env* e, const float* actions, long env0, long n_env, int first, float* obs, float* share_obs,
                        float* ava, float* reward, uint8_t* done, float* delay, float* payment, void* stream);

/* Preset contexts (Env(preset=True) / modify_preset, …SingleProcess.py:25-32,174-194,344-353):
 * overwrite the context the next step will run against and re-emit the observation.
 * R, C [B] float64; worker_pr [B*W] float64; n_disabled [B] int32; any may be NULL (= keep). */
int dcml_env_set_context(dcml_env* e, const double* R, const double* C, const double* worker_pr,
                         const int32_t* n_disabled, float* obs, float* share_obs, float* ava, void* stream);

/* ------------------------------------------------------------------------------------------
 * GAE / returns: replaces SharedReplayBuffer.compute_returns + ValueNorm.denormalize
 * (mat_src/mat/utils/shared_buffer.py:207-238, mat_src/mat/utils/valuenorm.py:32-36,69-79).
 *   rewards [T][n], value_preds [T+1][n] (row T = bootstrap value, already written),
 *   masks [T+1][n], n = N*L;  vn_stats = {running_mean, running_mean_sq, debiasing_term}
 *   device fp32[3] or NULL (no ValueNorm) -> advantages [T][n], returns [T+1][n] (row T untouched).
 * ------------------------------------------------------------------------------------------ */
int matas_gae(const float* rewards, const float* value_preds, const float* masks, const float* vn_stats,
              double gamma, double gae_lambda, int T, long n, float* advantages, float* returns, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused gradient clipping + Adam on the flat parameter vector: replaces
 * nn.utils.clip_grad_norm_(policy.transformer.parameters(), max_grad_norm) followed by
 * policy.optimizer.step() in MATTrainer.ppo_update (mat_src/mat/algorithms/mat/mat_trainer.py:149-154)
 * for optimizer = torch.optim.Adam(lr, eps, weight_decay) (.../transformer_policy.py:103-105).
 *   flat_params, flat_grad, exp_avg, exp_avg_sq  [n] fp32, 16-byte aligned; all four are updated in
 *   place (the gradient is scaled by the clip coefficient, as clip_grad_norm_ does)
 *   max_grad_norm <= 0: no clipping (get_gard_norm branch, mat_trainer.py:152); step: 1-based count
 *   norm_out  device float[2]: {total gradient norm before clipping, applied coefficient}
 * Call matas_policy_set_weights afterwards.
 * ------------------------------------------------------------------------------------------ */
int matas_adam_clip_step(float* flat_params, float* flat_grad, float* exp_avg, float* exp_avg_sq, long n,
                         double lr, double beta1, double beta2, double eps, double weight_decay,
                         double max_grad_norm, long step, float* norm_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MATAS_B200_H */
