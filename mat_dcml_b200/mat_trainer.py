"""Drop-in for the reference's ``MATTrainer`` (mat_src/mat/algorithms/mat/mat_trainer.py:11-223):
same constructor, ``train(buffer)`` / ``ppo_update(sample)`` / ``cal_value_loss`` / ``prep_*``.

The model forward/backward are the CUDA kernels behind ``policy.evaluate_actions``; the PPO
surrogate, the clipped Huber value loss and the ValueNorm update are the reference's few
element-wise expressions over (M*L, 1) device vectors.  Gradients accumulate in the model's flat
buffer; with ``torch.distributed`` initialised they are all-reduced (NCCL) once per minibatch, the
only collective on the path, together with the scalars the reference computes over
the global batch (advantage moments, ValueNorm batch moments, active-mask denominators).
"""
import numpy as np
import torch
import torch.distributed as dist
import torch.nn as nn

from .valuenorm import ValueNorm


def _world():
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def _allsum(t):
    if _world() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t


def huber_loss(e, d):                       # mat/utils/util.py:23-26
    a = (abs(e) <= d).float()
    b = (e > d).float()
    return a * e ** 2 / 2 + b * d * (abs(e) - d / 2)


def mse_loss(e):
    return e ** 2 / 2


class MATTrainer:
    def __init__(self, args, policy, num_agents, device=torch.device("cpu")):
        self.device = torch.device(device)
        self.tpdv = dict(dtype=torch.float32, device=self.device)
        self.policy, self.num_agents = policy, num_agents
        self.n_rollout_threads = args.n_rollout_threads
        self.clip_param, self.ppo_epoch, self.num_mini_batch = args.clip_param, args.ppo_epoch, args.num_mini_batch
        self.value_loss_coef, self.entropy_coef = args.value_loss_coef, args.entropy_coef
        self.max_grad_norm, self.huber_delta = args.max_grad_norm, args.huber_delta
        self._use_max_grad_norm = args.use_max_grad_norm
        self._use_clipped_value_loss = args.use_clipped_value_loss
        self._use_huber_loss = args.use_huber_loss
        self._use_valuenorm = args.use_valuenorm
        self._use_value_active_masks = args.use_value_active_masks
        self._use_policy_active_masks = args.use_policy_active_masks
        self.value_normalizer = ValueNorm(1, device=self.device) if self._use_valuenorm else None

    def _t(self, x):
        if isinstance(x, np.ndarray):
            x = torch.from_numpy(x)
        return x.to(**self.tpdv)

    def _vn_update(self, returns, moments=None):
        """ValueNorm.update over the GLOBAL minibatch (valuenorm.py:39-56).  ``moments`` = (sum, sum of squares, count) of the
        global minibatch's returns when the caller has already reduced them (train() does it once per epoch for all
        minibatches: they depend on the returns and the permutation only)."""
        if _world() == 1:
            self.value_normalizer.update(returns)
            return
        if moments is not None:
            m = moments
        else:
            m = torch.stack([returns.sum(), (returns ** 2).sum(), torch.tensor(float(returns.numel()), **self.tpdv)])
            _allsum(m)
        vn, w = self.value_normalizer, self.value_normalizer.beta
        vn.stats[0:1].mul_(w).add_((m[0] / m[2]) * (1.0 - w))
        vn.stats[1:2].mul_(w).add_((m[1] / m[2]) * (1.0 - w))
        vn.stats[2].mul_(w).add_(1.0 * (1.0 - w))

    def cal_value_loss(self, values, value_preds_batch, return_batch, active_masks_batch, vn_moments=None, numerator_only=False):
        """mat_trainer.py:54-94.  ``numerator_only``: the masked SUM of the loss without the division by the active count
        (ppo_update divides the all-reduced gradient once instead, see there)."""
        value_pred_clipped = value_preds_batch + (values - value_preds_batch).clamp(-self.clip_param, self.clip_param)
        if self._use_valuenorm:
            self._vn_update(return_batch, vn_moments)
            error_clipped = self.value_normalizer.normalize(return_batch) - value_pred_clipped
            error_original = self.value_normalizer.normalize(return_batch) - values
        else:
            error_clipped = return_batch - value_pred_clipped
            error_original = return_batch - values
        if self._use_huber_loss:
            vl_clipped, vl_original = huber_loss(error_clipped, self.huber_delta), huber_loss(error_original, self.huber_delta)
        else:
            vl_clipped, vl_original = mse_loss(error_clipped), mse_loss(error_original)
        value_loss = torch.max(vl_original, vl_clipped) if self._use_clipped_value_loss else vl_original
        if numerator_only:
            return (value_loss * active_masks_batch).sum() if self._use_value_active_masks else value_loss.sum()
        if self._use_value_active_masks:
            return (value_loss * active_masks_batch).sum() / self._active_sum(active_masks_batch)
        return value_loss.sum() / self._count(value_loss)

    def _active_sum(self, am):
        return _allsum(am.sum().detach().clone())

    def _count(self, x):
        return _allsum(torch.tensor(float(x.numel()), **self.tpdv))

    def ppo_update(self, sample, vn_moments=None):
        """mat_trainer.py:96-156.  Under torch.distributed every rank passes its local minibatch and all replicas take the
        step the reference would take on the concatenated batch, with ONE collective: the three loss terms share one
        denominator (the global active count, or the global element count), so each rank back-propagates the un-normalised
        numerators, the flat gradient travels with the local count and the three numerators appended, and the sum is
        divided once (``vn_moments``: the ValueNorm batch moments of the global minibatch, reduced per epoch by train())."""
        share_obs_batch, obs_batch, rnn_states_batch, rnn_states_critic_batch, actions_batch, \
            value_preds_batch, return_batch, masks_batch, active_masks_batch, old_action_log_probs_batch, \
            adv_targ, available_actions_batch = sample
        old_action_log_probs_batch = self._t(old_action_log_probs_batch)
        adv_targ, value_preds_batch = self._t(adv_targ), self._t(value_preds_batch)
        return_batch, active_masks_batch = self._t(return_batch), self._t(active_masks_batch)
        model = self.policy.transformer
        model.flat_grad(zero=True)                                   # optimizer.zero_grad()
        fused = _world() > 1 and self._use_policy_active_masks == self._use_value_active_masks
        one = torch.ones((), **self.tpdv)
        den = one if fused else self._active_sum(active_masks_batch)
        values, action_log_probs, dist_entropy = self.policy.evaluate_actions(
            share_obs_batch, obs_batch, rnn_states_batch, rnn_states_critic_batch, actions_batch, masks_batch,
            available_actions_batch, active_masks_batch, active_denominator=den)
        if fused and not self._use_policy_active_masks:
            dist_entropy = dist_entropy * float(action_log_probs.shape[0])       # .mean() back to a sum
        if values.is_cuda:
            policy_loss, value_loss, imp_weights, cnt = self._loss_backward_cuda(
                values, action_log_probs, dist_entropy, old_action_log_probs_batch, adv_targ, value_preds_batch, return_batch,
                active_masks_batch, den, fused, vn_moments)
        else:
            # host tensors (the gloo tests of the sharded update): the reference's expressions, walked back by autograd
            imp_weights = torch.exp(action_log_probs - old_action_log_probs_batch)
            surr1 = imp_weights * adv_targ
            surr2 = torch.clamp(imp_weights, 1.0 - self.clip_param, 1.0 + self.clip_param) * adv_targ
            if self._use_policy_active_masks:
                policy_loss = (-torch.sum(torch.min(surr1, surr2) * active_masks_batch, dim=-1, keepdim=True)).sum() / den
            else:
                policy_loss = -torch.sum(torch.min(surr1, surr2), dim=-1, keepdim=True).sum() / (one if fused else self._count(surr1))
            value_loss = self.cal_value_loss(values, value_preds_batch, return_batch, active_masks_batch, vn_moments, numerator_only=fused)
            loss = policy_loss - dist_entropy * self.entropy_coef + value_loss * self.value_loss_coef
            loss.backward()                                          # CUDA backward -> flat gradient
            cnt = active_masks_batch.sum() if self._use_policy_active_masks else torch.tensor(float(surr1.numel()), **self.tpdv)
        if fused:
            # the one collective of the update: [flat gradient | local denominator, policy / value / entropy numerators]
            comm = model.flat_grad_comm()
            comm[-4:] = torch.stack([cnt.detach(), policy_loss.detach(), value_loss.detach(), dist_entropy.detach()])
            dist.all_reduce(comm, op=dist.ReduceOp.SUM)
            inv = 1.0 / comm[-4]
            comm[:-4].mul_(inv)                                      # = model.flat_grad(): gradient of the global-batch loss
            policy_loss, value_loss, dist_entropy = comm[-3] * inv, comm[-2] * inv, comm[-1] * inv
        elif _world() > 1:
            dist.all_reduce(model.flat_grad(), op=dist.ReduceOp.SUM)
        opt = self.policy.optimizer
        if hasattr(opt, "clip_and_step"):                            # FlatAdam: clip + Adam in two launches
            grad_norm = opt.clip_and_step(self.max_grad_norm if self._use_max_grad_norm else 0.0)
        else:
            params = [p for p in model.parameters() if p.grad is not None]
            if self._use_max_grad_norm:
                grad_norm = nn.utils.clip_grad_norm_(params, self.max_grad_norm)
            else:
                grad_norm = torch.sqrt(sum((p.grad ** 2).sum() for p in params))
            opt.step()
        return value_loss, grad_norm, policy_loss, dist_entropy, grad_norm, imp_weights

    def _loss_backward_cuda(self, values, action_log_probs, dist_entropy, old_logp, adv, value_preds, returns, active, den, fused,
                            vn_moments):
        """The surrogate, the value loss and autograd's pass through them as ONE launch (``matas_ppo_loss``,
        mat_trainer.py:54-94,121-137), then the model's backward from the three gradients.  Returns
        (policy_loss, value_loss, importance weights, local denominator); with ``fused`` (sharded minibatch) the two losses are
        the un-normalised numerators, as in ppo_update's collective."""
        from . import _cabi
        n = values.numel()
        c = lambda t: t.contiguous().view(-1)
        if self._use_valuenorm:
            self._vn_update(returns, vn_moments)                     # cal_value_loss updates the running statistics first
            mean, var = self.value_normalizer.running_mean_var()
            vn = torch.cat([mean.view(1), torch.sqrt(var).view(1)]).to(**self.tpdv)
        else:
            vn = None
        count = torch.tensor(float(n), **self.tpdv)
        cnt = active.sum() if self._use_policy_active_masks else count
        if fused:
            dens = torch.ones(2, **self.tpdv)
        else:
            den_p = den if self._use_policy_active_masks else self._count(values)
            den_v = den if self._use_value_active_masks else self._count(values)
            dens = torch.stack([den_p.reshape(()), den_v.reshape(())]).to(**self.tpdv)
        g_lp, g_v = torch.empty_like(action_log_probs), torch.empty_like(values)
        ratio = torch.empty_like(action_log_probs)
        sums = torch.empty(2, dtype=torch.float64, device=values.device)
        lp, vv = action_log_probs.detach(), values.detach()
        assert lp.is_contiguous() and vv.is_contiguous()
        a_old, a_adv, a_act, a_vp, a_ret = c(old_logp), c(adv), c(active), c(value_preds), c(returns)
        _cabi.check(_cabi.lib().matas_ppo_loss(
            n, _cabi.ptr(lp), _cabi.ptr(a_old), _cabi.ptr(a_adv), _cabi.ptr(a_act), _cabi.ptr(vv), _cabi.ptr(a_vp), _cabi.ptr(a_ret),
            _cabi.ptr(vn) if vn is not None else None, _cabi.ptr(dens), float(self.clip_param), float(self.huber_delta),
            float(self.value_loss_coef), int(self._use_huber_loss), int(self._use_clipped_value_loss),
            int(self._use_policy_active_masks), int(self._use_value_active_masks), _cabi.ptr(g_lp), _cabi.ptr(g_v), _cabi.ptr(ratio),
            _cabi.ptr(sums), _cabi.stream_ptr()))
        policy_loss = (-sums[0] / dens[0]).float()
        value_loss = (sums[1] / dens[1]).float()
        g_ent = torch.full_like(dist_entropy, -self.entropy_coef)
        torch.autograd.backward([action_log_probs, values, dist_entropy], [g_lp, g_v, g_ent])   # CUDA backward -> flat gradient
        return policy_loss, value_loss, ratio, cnt

    def normalize_advantages(self, adv, active):
        """mat_trainer.py:193-197: (adv - nanmean) / (nanstd + 1e-5) over active entries of the GLOBAL batch; one collective
        (count, sum, sum of squares in float64) when the batch is sharded."""
        if _world() == 1:
            m = torch.stack([(adv * active).sum(), active.sum()]).double()
            mean = m[0] / m[1]
            std = torch.sqrt((((adv.double() - mean) ** 2) * active).sum() / m[1])
            return ((adv - mean.float()) / (std.float() + 1e-5))
        a = adv.double()
        m = torch.stack([(a * active).sum(), (a * a * active).sum(), active.sum().double()])
        _allsum(m)
        mean = m[0] / m[2]
        std = torch.sqrt((m[1] / m[2] - mean * mean).clamp(min=0.0))
        return ((adv - mean.float()) / (std.float() + 1e-5))

    def minibatch_moments(self, samples):
        """ValueNorm batch moments (sum, sum of squares, count of the returns, valuenorm.py:45-46) of every minibatch of an
        epoch over the GLOBAL batch: one collective per epoch instead of one per minibatch."""
        if _world() == 1 or not self._use_valuenorm:
            return [None] * len(samples)
        m = torch.stack([torch.stack([self._t(s[6]).sum(), (self._t(s[6]) ** 2).sum(),
                                      torch.tensor(float(self._t(s[6]).numel()), **self.tpdv)]) for s in samples])
        _allsum(m)
        return list(m)

    def train(self, buffer):
        """mat_trainer.py:158-217"""
        # the statistics are accumulated on the device and read back once: a .item() per minibatch (as in the
        # reference, mat_trainer.py:199-205) would stall the host after every update
        acc = torch.zeros(6, dtype=torch.float64, device=self.device)
        for _ in range(self.ppo_epoch):
            next_values = self.policy.get_values(None, buffer.obs[-1].reshape(-1, buffer.obs.shape[-1]), None, None)
            buffer.compute_returns(next_values.reshape(buffer.value_preds[-1].shape), self.value_normalizer)
            advantages = self.normalize_advantages(buffer.advantages, buffer.active_masks[:-1])
            samples = buffer.feed_forward_generator_transformer(advantages, self.num_mini_batch)
            if _world() > 1:
                samples = list(samples)                              # the epoch's minibatches: their ValueNorm moments in one collective
                moments = self.minibatch_moments(samples)
            for k, sample in enumerate(samples):
                vl, cgn, pl, ent, agn, ratio = self.ppo_update(sample, moments[k] if _world() > 1 else None)
                with torch.no_grad():
                    acc += torch.stack([vl.detach().double(), pl.detach().double(), ent.detach().double(),
                                        torch.as_tensor(agn, device=self.device).double(), torch.as_tensor(cgn, device=self.device).double(),
                                        ratio.detach().mean().double()])
        keys = ("value_loss", "policy_loss", "dist_entropy", "actor_grad_norm", "critic_grad_norm", "ratio")
        info = dict(zip(keys, acc.tolist()))
        n = self.ppo_epoch * self.num_mini_batch
        if _world() > 1 and self._use_policy_active_masks != self._use_value_active_masks:
            # (mixed denominators take the per-term collectives: each rank logged its share of the global losses)
            t = torch.tensor([info["value_loss"], info["policy_loss"], info["dist_entropy"]], dtype=torch.float64, device=self.device)
            _allsum(t)
            info["value_loss"], info["policy_loss"], info["dist_entropy"] = t.tolist()
        return {k: v / n for k, v in info.items()}

    def prep_training(self):
        self.policy.train()

    def prep_rollout(self):
        self.policy.eval()
