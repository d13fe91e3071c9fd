"""CPU restatement of GAE / returns, ValueNorm and the PPO loss (test infrastructure;
see oracle/__init__.py).  Citations: ``/root/reference/mat_src/mat`` files
``utils/shared_buffer.py`` (BUF), ``utils/valuenorm.py`` (VN),
``algorithms/mat/mat_trainer.py`` (TR), ``utils/util.py`` (UTIL).
"""
import numpy as np
import torch

from . import mat_oracle as mo


class ValueNormState:
    """VN:8-80 with ``input_shape=1``: three fp32 scalars (running mean, running mean of
    squares, debiasing term), beta = 0.99999."""

    def __init__(self, beta=0.99999, epsilon=1e-5):
        self.beta, self.epsilon = beta, epsilon
        self.running_mean = torch.zeros(1)
        self.running_mean_sq = torch.zeros(1)
        self.debiasing_term = torch.zeros(())

    def mean_var(self):                                                  # VN:32-36
        d = self.debiasing_term.clamp(min=self.epsilon)
        mean = self.running_mean / d
        var = (self.running_mean_sq / d - mean ** 2).clamp(min=1e-2)
        return mean, var

    @torch.no_grad()
    def update(self, x):                                                 # VN:39-56
        x = torch.as_tensor(x, dtype=torch.float32)
        bm, bsq = x.mean(dim=0), (x ** 2).mean(dim=0)
        w = self.beta
        self.running_mean.mul_(w).add_(bm * (1.0 - w))
        self.running_mean_sq.mul_(w).add_(bsq * (1.0 - w))
        self.debiasing_term.mul_(w).add_(1.0 * (1.0 - w))

    def normalize(self, x):                                              # VN:58-66
        mean, var = self.mean_var()
        return (torch.as_tensor(x, dtype=torch.float32) - mean) / torch.sqrt(var)

    def denormalize(self, x):                                            # VN:69-79
        mean, var = self.mean_var()
        return (torch.as_tensor(x, dtype=torch.float32) * torch.sqrt(var) + mean).numpy()

    def stats(self):
        return np.array([self.running_mean.item(), self.running_mean_sq.item(), self.debiasing_term.item()],
                        dtype=np.float32)


def compute_returns(rewards, value_preds, masks, next_value, vn=None, gamma=0.99, lam=0.95):
    """BUF:207-238 (GAE branch).  rewards (T,N,L,1); value_preds, masks (T+1,N,L,1) fp32.
    -> advantages (T,N,L,1), returns (T+1,N,L,1), value_preds with [-1] = next_value."""
    rewards = np.asarray(rewards, np.float32)
    v = np.array(value_preds, np.float32)
    masks = np.asarray(masks, np.float32)
    v[-1] = next_value
    T = rewards.shape[0]
    adv = np.zeros_like(rewards)
    ret = np.zeros_like(v)
    den = (lambda x: x) if vn is None else vn.denormalize
    gae = 0
    for t in reversed(range(T)):
        delta = rewards[t] + gamma * den(v[t + 1]) * masks[t + 1] - den(v[t])
        gae = delta + gamma * lam * masks[t + 1] * gae
        adv[t] = gae
        ret[t] = gae + den(v[t])
    return adv, ret, v


def normalize_advantages(adv, active_masks):
    """TR:193-197: nan-masked mean/std over active entries, eps 1e-5."""
    a = np.array(adv, copy=True)
    a[np.asarray(active_masks)[:-1] == 0.0] = np.nan
    return (adv - np.nanmean(a)) / (np.nanstd(a) + 1e-5)


def huber(e, d):                                                         # UTIL:23-26
    a = (e.abs() <= d).float()
    b = (e > d).float()
    return a * e ** 2 / 2 + b * d * (e.abs() - d / 2)


def ppo_loss(sd, n_agent, obs, actions, ava, old_logp, adv, value_preds, returns, active, vn,
             clip=0.2, entropy_coef=0.01, value_loss_coef=1.0, huber_delta=10.0, mode="semi"):
    """TR:96-144 up to ``loss``; updates ``vn`` like cal_value_loss does (TR:69).
    -> (loss, value_loss, policy_loss, entropy, ratio)."""
    f = lambda x: torch.as_tensor(np.asarray(x) if not torch.is_tensor(x) else x, dtype=torch.float32).reshape(-1, 1)
    old_logp, adv, value_preds, returns, active = map(f, (old_logp, adv, value_preds, returns, active))
    values, logp, ent = mo.evaluate_actions(sd, obs, actions, ava, n_agent, active, mode)
    ratio = torch.exp(logp - old_logp)
    s1 = ratio * adv
    s2 = torch.clamp(ratio, 1.0 - clip, 1.0 + clip) * adv
    pol = (-torch.sum(torch.min(s1, s2) * active, dim=-1, keepdim=True)).sum() / active.sum()
    vclip = value_preds + (values - value_preds).clamp(-clip, clip)
    vn.update(returns)
    e_clip = vn.normalize(returns) - vclip
    e_orig = vn.normalize(returns) - values
    vl = torch.max(huber(e_orig, huber_delta), huber(e_clip, huber_delta))
    vl = (vl * active).sum() / active.sum()
    loss = pol - ent * entropy_coef + vl * value_loss_coef
    return loss, vl, pol, ent, ratio
