"""GPU parity of the evaluate backward / PPO update against torch autograd on the CPU oracle and
against the reference's own MATTrainer.ppo_update (golden fixture)."""
import argparse

import numpy as np
import pytest
import torch

from conftest import golden
from oracle import mat_oracle as mo
from oracle import ppo_oracle as po
from test_policy_gpu import make_model

pytestmark = pytest.mark.gpu


def _targs(**over):
    ns = argparse.Namespace(algorithm_name="mat", env_name="DCML", lr=0.0, opti_eps=1e-5, weight_decay=0,
                            use_policy_active_masks=True, n_block=2, n_embd=64, n_head=2, encode_state=False,
                            dec_actor=False, share_actor=False, n_rollout_threads=1, clip_param=0.2, ppo_epoch=15,
                            num_mini_batch=4, data_chunk_length=10, value_loss_coef=1.0, entropy_coef=0.01,
                            max_grad_norm=1e9, huber_delta=10.0, use_recurrent_policy=False,
                            use_naive_recurrent_policy=False, use_max_grad_norm=True, use_clipped_value_loss=True,
                            use_huber_loss=True, use_valuenorm=True, use_value_active_masks=True)
    for k, v in over.items():
        setattr(ns, k, v)
    return ns


def _grad_check(model, sd_grads, rel=2e-3):
    """per-tensor: |g - g_ref| <= rel * max(|g_ref|_2-norm scale, floor)"""
    total = np.sqrt(sum(float((g.double() ** 2).sum()) for g in sd_grads.values() if g is not None))
    worst = 0.0
    for name, p in model.named_parameters():
        want = sd_grads[name]
        got = p.grad.detach().cpu()
        if want is None:
            assert float(got.abs().max()) == 0.0, name + " should have no gradient"
            continue
        scale = max(float(want.norm()), 1e-4 * total)
        err = float((got - want).norm()) / scale
        worst = max(worst, err)
        assert err < rel, "%s: relative gradient error %.3e" % (name, err)
    return worst


@pytest.mark.parametrize("L,M,semi", [(11, 6, True), (21, 13, True), (9, 70, False), (3, 1, True), (51, 5, True), (64, 3, True),
                                          (65, 2, True), (64, 2, False)])
def test_evaluate_backward_vs_autograd(L, M, semi, cuda_device):
    sd = mo.make_weights(50 + L, L, semi=semi)
    for v in sd.values():
        v.requires_grad_(True)
    rng = np.random.RandomState(L)
    obs = rng.uniform(0, 1.1, size=(M * L, 7)).astype(np.float32)
    ava = np.ones((M, L, 2), np.float32)
    ava[rng.uniform(size=(M, L)) < 0.3, 1] = 0
    if semi:
        ava[:, -1, 1] = 1
    act = (rng.uniform(size=(M, L, 1)) < 0.5).astype(np.float32) * ava[:, :, 1:2]
    if semi:
        act[:, -1, 0] = rng.standard_normal(M)
    gv, glp, gent = (rng.standard_normal((M * L, 1)).astype(np.float32) for _ in range(3))
    mode = "semi" if semi else "discrete"
    # oracle: torch autograd on CPU
    obs_t = torch.from_numpy(obs).reshape(M, L, 7)
    v, rep = mo.encoder(sd, obs_t)
    fn = mo.semi_discrete_parallel_act if semi else mo.discrete_parallel_act
    lp, ent = fn(sd, rep, torch.from_numpy(act), torch.from_numpy(ava))
    loss = (v.reshape(-1, 1) * torch.from_numpy(gv)).sum() + (lp.reshape(-1, 1) * torch.from_numpy(glp)).sum() + \
        (ent.reshape(-1, 1) * torch.from_numpy(gent)).sum()
    loss.backward()
    want = {k: t.grad for k, t in sd.items()}
    # kernels
    m = make_model(L, {k: t.detach() for k, t in sd.items()}, semi=semi)
    m.flat_grad(zero=True)
    lp2, v2, ent2 = m(None, obs, act, ava)
    np.testing.assert_allclose(lp2.detach().cpu().numpy().reshape(-1), lp.detach().numpy().reshape(-1), atol=1e-4)
    loss2 = (v2.reshape(-1, 1) * torch.from_numpy(gv).cuda()).sum() + (lp2.reshape(-1, 1) * torch.from_numpy(glp).cuda()).sum() + \
        (ent2.reshape(-1, 1) * torch.from_numpy(gent).cuda()).sum()
    loss2.backward()
    _grad_check(m, want)
    # accumulation semantics: a second backward doubles the gradient
    g1 = m.flat_grad().clone()
    lp3, v3, ent3 = m(None, obs, act, ava)
    ((v3.reshape(-1, 1) * torch.from_numpy(gv).cuda()).sum() + (lp3.reshape(-1, 1) * torch.from_numpy(glp).cuda()).sum() +
     (ent3.reshape(-1, 1) * torch.from_numpy(gent).cuda()).sum()).backward()
    assert float((m.flat_grad() - 2 * g1).abs().max()) <= 1e-3 * float(g1.abs().max())


def test_ppo_update_vs_reference(cuda_device):
    """One MATTrainer.ppo_update: losses and per-tensor gradient statistics of the reference."""
    from mat_dcml_b200 import TransformerPolicy, Action_Space, MATTrainer
    g = golden("ppo_L11")
    L, M = int(g["L"]), int(g["M"])
    args = _targs()
    pol = TransformerPolicy(args, [7], [L + 1], Action_Space(2, semi_index=-1, extra=True), L, device=torch.device("cuda"))
    pol.transformer.load_state_dict(mo.full_state_dict(mo.make_weights(int(g["wseed"]), L), L))
    tr = MATTrainer(args, pol, L, device=torch.device("cuda"))
    tr.value_normalizer.stats.copy_(torch.from_numpy(g["vn_stats_before"]))
    sample = (None, g["obs"].astype(np.float32), None, None, g["actions"], g["value_preds"], g["returns"], None,
              g["active"], g["old_logp"], g["adv"], g["ava"].astype(np.float32))
    vl, gn, pl, ent, _, ratio = tr.ppo_update(sample)
    assert abs(vl.item() - float(g["value_loss"])) < 1e-4 * max(1.0, abs(float(g["value_loss"])))
    assert abs(pl.item() - float(g["policy_loss"])) < 1e-4
    assert abs(ent.item() - float(g["entropy"])) < 1e-4
    assert abs(ratio.mean().item() - float(g["ratio_mean"])) < 1e-4
    assert abs(float(gn) - float(g["grad_norm"])) < 2e-3 * float(g["grad_norm"])
    proj = np.random.RandomState(99)
    params = dict(pol.transformer.named_parameters())
    for name, st in zip(g["grad_names"], g["grad_stats"]):
        gr = params[str(name)].grad.detach().cpu().numpy()
        r = proj.standard_normal(gr.shape)
        scale = max(st[1], 1e-4 * float(g["grad_norm"]))
        assert abs(np.sqrt((gr.astype(np.float64) ** 2).sum()) - st[1]) < 3e-3 * scale, name
        assert abs((gr * r).sum(dtype=np.float64) - st[2]) < 3e-3 * scale * np.sqrt(gr.size), name
        if not st[3]:
            assert np.abs(gr).max() == 0.0, name
        if "grad/" + str(name) in g.files:
            np.testing.assert_allclose(gr, g["grad/" + str(name)], atol=3e-3 * scale)


def test_training_step_changes_weights_and_full_loop(cuda_device):
    """rollout -> GAE -> 2 PPO epochs on a small env batch: finite, weights move, value loss drops."""
    from mat_dcml_b200 import TransformerPolicy, MATTrainer, DcmlVecEnv, SharedReplayBuffer
    W, L, N, T = 10, 11, 32, 8
    args = _targs(lr=5e-4, max_grad_norm=10.0, ppo_epoch=2, num_mini_batch=2, episode_length=T, n_rollout_threads=N,
                  hidden_size=64, recurrent_N=1, gamma=0.99, gae_lambda=0.95, use_gae=True, use_popart=True,
                  use_proper_time_limits=False)
    env = DcmlVecEnv(N, W, seed=3)
    pol = TransformerPolicy(args, [7], [W + 2], env.action_space[0], L, device=torch.device("cuda"))
    pol.transformer.output_device = torch.device("cuda")
    tr = MATTrainer(args, pol, L, device=torch.device("cuda"))
    buf = SharedReplayBuffer(args, L, [7], [W + 2], env.action_space[0], "DCML")
    obs, sob, ava = env.reset_device()
    buf.obs[0].copy_(obs); buf._share_obs[0].copy_(env.sob); buf.available_actions[0].copy_(ava)
    for t in range(T):
        v, a, lp, _, _ = pol.get_actions(None, obs.reshape(-1, 7), None, None, None, ava.reshape(-1, 2))
        obs, sob, rew, done, delay, pay, ava = env.step_device(a.reshape(N, L))
        buf.insert(env.sob, obs, None, None, a.reshape(N, L, 1), lp.reshape(N, L, 1), v.reshape(N, L, 1), rew,
                   1.0 - done.float(), None, None, ava)
    w0 = pol.transformer.flat_parameters().clone()
    info = tr.train(buf)
    buf.after_update()
    assert all(np.isfinite(v) for v in info.values()), info
    assert float((pol.transformer.flat_parameters() - w0).abs().max()) > 0
    assert 0.5 < info["ratio"] < 1.5


@pytest.mark.parametrize("max_norm,wd", [(0.5, 0.0), (1e9, 0.0), (0.0, 0.0), (10.0, 1e-3)])
def test_fused_clip_adam_vs_torch(max_norm, wd, cuda_device):
    """matas_adam_clip_step (FlatAdam) against nn.utils.clip_grad_norm_ + torch.optim.Adam on per-tensor copies
    (the reference's update, mat_trainer.py:149-154 / transformer_policy.py:103-105) over three steps."""
    from mat_dcml_b200.flat_adam import FlatAdam
    L = 11
    m = make_model(L, mo.make_weights(7, L))
    opt = FlatAdam(m, lr=3e-3, eps=1e-5, weight_decay=wd)
    ref = [torch.nn.Parameter(p.detach().clone()) for p in m.parameters()]
    ropt = torch.optim.Adam(ref, lr=3e-3, eps=1e-5, weight_decay=wd)
    gen = torch.Generator(device="cuda").manual_seed(5)
    # encoder.state_encoder.* / decoder.obs_encoder.* never receive a gradient (ma_transformer.py:144-149,198-199): in the
    # reference their .grad stays None and torch.optim.Adam skips them, weight decay included
    frozen = [n.startswith("encoder.state_encoder.") or n.startswith("decoder.obs_encoder.") for n, _ in m.named_parameters()]
    for it in range(3):
        g = torch.randn(m.flat_parameters().numel(), device="cuda", generator=gen) * (0.02 if it else 2.0)
        off = 0
        for p, fz in zip(ref, frozen):
            if fz:
                g[off:off + p.numel()] = 0
                p.grad = None
            else:
                p.grad = g[off:off + p.numel()].view(p.shape).clone()
            off += p.numel()
        m.flat_grad(zero=True).copy_(g)
        if max_norm > 0:
            want_norm = torch.nn.utils.clip_grad_norm_(ref, max_norm)
        else:
            want_norm = torch.sqrt(sum((p.grad ** 2).sum() for p in ref if p.grad is not None))
        ropt.step()
        ver = m.flat_parameters()._version
        got_norm = opt.clip_and_step(max_norm)
        assert abs(float(got_norm) - float(want_norm)) <= 1e-5 * float(want_norm)
        want = torch.cat([p.detach().reshape(-1) for p in ref])
        np.testing.assert_allclose(m.flat_parameters().cpu().numpy(), want.cpu().numpy(), rtol=2e-5, atol=2e-7)
        want_g = torch.cat([(torch.zeros_like(p) if p.grad is None else p.grad).reshape(-1) for p in ref])
        np.testing.assert_allclose(m.flat_grad().cpu().numpy(), want_g.cpu().numpy(), rtol=1e-5, atol=1e-9)
        assert m._packed_key is None                      # the next pass repacks the weights
    assert opt.state_dict()["step"] == 3
    # the policy picks it up on a GPU, and lr_decay reaches it through param_groups
    from mat_dcml_b200 import TransformerPolicy, Action_Space
    pol = TransformerPolicy(_targs(lr=5e-4), [7], [L + 1], Action_Space(2, semi_index=-1, extra=True), L, device=torch.device("cuda"))
    assert isinstance(pol.optimizer, FlatAdam)
    pol.lr_decay(1, 2)
    assert abs(pol.optimizer.param_groups[0]["lr"] - 2.5e-4) < 1e-12


@pytest.mark.parametrize("flags", [(1, 1, 1, 1), (0, 1, 1, 1), (1, 0, 0, 1), (1, 1, 0, 0), (0, 0, 1, 0)])
@pytest.mark.parametrize("valuenorm", [True, False])
def test_ppo_loss_kernel_vs_torch_autograd(flags, valuenorm, cuda_device):
    """matas_ppo_loss (one launch: surrogate + value loss + their gradients) against the reference's element-wise expressions
    (mat_trainer.py:54-94,121-137; util.py:23-29) walked back by torch autograd, on inputs that hit every branch: ratios inside /
    outside / exactly on the clip bounds, exact min / max ties, errors beyond +-huber_delta, dead (masked) agents."""
    import ctypes
    from mat_dcml_b200 import _cabi
    from mat_dcml_b200.mat_trainer import huber_loss, mse_loss
    use_huber, use_clipped, pol_mask, val_mask = flags
    dev = cuda_device
    g = torch.Generator(device="cpu").manual_seed(5)
    n = 40_000
    clip, delta, vcoef = 0.2, 10.0, 0.7
    old = torch.randn(n, 1, generator=g) * 0.5 - 1.0
    lp = old + torch.randn(n, 1, generator=g) * 0.25
    lp[:500] = old[:500]                                            # ratio exactly 1: min tie inside the clip range
    lp[500:600] = old[500:600] + float(np.log(1.2))                  # (nearly) on the upper bound
    adv = torch.randn(n, 1, generator=g)
    adv[600:700] = 0.0                                              # both surrogates zero
    vp = torch.randn(n, 1, generator=g) * 3
    v = vp + torch.randn(n, 1, generator=g) * 0.3
    v[700:900] = vp[700:900]                                        # clipped == unclipped: max tie
    ret = torch.randn(n, 1, generator=g) * 40 - 100                  # normalised: O(1); raw: errors far beyond +-delta
    act = (torch.rand(n, 1, generator=g) > 0.1).float()
    lp, old, adv, vp, v, ret, act = [t.to(dev) for t in (lp, old, adv, vp, v, ret, act)]
    mean, std = torch.tensor([-98.0], device=dev), torch.tensor([37.5], device=dev)
    # ---- torch ----
    lp_t, v_t = lp.clone().requires_grad_(True), v.clone().requires_grad_(True)
    r = torch.exp(lp_t - old)
    m = torch.min(r * adv, torch.clamp(r, 1.0 - clip, 1.0 + clip) * adv)
    den_p = act.sum() if pol_mask else torch.tensor(float(n), device=dev)
    den_v = act.sum() if val_mask else torch.tensor(float(n), device=dev)
    pol_num = (m * act).sum() if pol_mask else m.sum()
    tgt = (ret - mean) / std if valuenorm else ret
    vc = vp + (v_t - vp).clamp(-clip, clip)
    L = (lambda e: huber_loss(e, delta)) if use_huber else mse_loss
    lo, lc = L(tgt - v_t), L(tgt - vc)
    vl = torch.max(lo, lc) if use_clipped else lo
    val_num = (vl * act).sum() if val_mask else vl.sum()
    (-pol_num / den_p + vcoef * val_num / den_v).backward()
    # ---- kernel ----
    g_lp, g_v, ratio = torch.empty_like(lp), torch.empty_like(v), torch.empty_like(lp)
    sums = torch.empty(2, dtype=torch.float64, device=dev)
    dens = torch.stack([den_p, den_v]).float()
    vn = torch.cat([mean, std]) if valuenorm else None
    P = _cabi.ptr
    _cabi.check(_cabi.lib().matas_ppo_loss(n, P(lp), P(old), P(adv), P(act), P(v), P(vp), P(ret), P(vn) if vn is not None else None, P(dens),
                                           clip, delta, vcoef, use_huber, use_clipped, pol_mask, val_mask, P(g_lp), P(g_v), P(ratio), P(sums),
                                           _cabi.stream_ptr()))
    torch.cuda.synchronize()
    assert torch.allclose(ratio, r.detach(), rtol=1e-6, atol=0)
    pol_num, val_num = pol_num.detach(), val_num.detach()
    assert abs(float(sums[0]) - float(pol_num)) <= 1e-5 * abs(float(pol_num)) + 1e-3
    assert abs(float(sums[1]) - float(val_num)) <= 1e-5 * abs(float(val_num)) + 1e-3
    scale_p, scale_v = float(lp_t.grad.abs().max()), float(v_t.grad.abs().max())
    assert float((g_lp - lp_t.grad).abs().max()) <= 2e-6 * scale_p + 1e-12
    assert float((g_v - v_t.grad).abs().max()) <= 2e-6 * scale_v + 1e-12
    # branches were hit
    rr = r.detach()
    assert int(((rr < 1 - clip) | (rr > 1 + clip)).sum()) > 100 and int((rr == 1.0).sum()) >= 500
    if not valuenorm and use_huber:
        assert int(((ret - v) > delta).sum()) > 0 and int(((ret - v) < -delta).sum()) > 100
