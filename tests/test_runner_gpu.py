"""Device-resident runner (mat_dcml_b200/dcml_runner.py) against a rollout composed from the oracle's pieces:
same named Philox draws -> same actions, log-probs, values, rewards, masks and GAE returns in the buffer as the
reference's DCMLRunner.collect / insert / compute would produce (dcml_runner.py:22-288, base_runner.py:286-325)."""
import argparse

import numpy as np
import pytest
import torch

from oracle import env_oracle as eo
from oracle import mat_oracle as mo
from oracle import philox as px
from oracle import ppo_oracle as po

pytestmark = pytest.mark.gpu


def _args(**over):
    ns = argparse.Namespace(algorithm_name="mat", env_name="DCML", lr=5e-4, opti_eps=1e-5, weight_decay=0,
                            use_policy_active_masks=True, n_block=2, n_embd=64, n_head=2, encode_state=False,
                            dec_actor=False, share_actor=False, clip_param=0.2, ppo_epoch=2, num_mini_batch=2,
                            data_chunk_length=10, value_loss_coef=1.0, entropy_coef=0.01, max_grad_norm=10.0, huber_delta=10.0,
                            use_recurrent_policy=False, use_naive_recurrent_policy=False, use_max_grad_norm=True,
                            use_clipped_value_loss=True, use_huber_loss=True, use_valuenorm=True, use_value_active_masks=True,
                            hidden_size=64, recurrent_N=1, gamma=0.99, gae_lambda=0.95, use_gae=True, use_popart=True,
                            use_proper_time_limits=False, use_linear_lr_decay=True, use_centralized_V=True, log_interval=1,
                            num_env_steps=0, episode_length=6, n_rollout_threads=8)
    for k, v in over.items():
        setattr(ns, k, v)
    return ns


def test_runner_rollout_matches_oracle_composition(cuda_device):
    from mat_dcml_b200 import DcmlRunner, DcmlVecEnv
    W, L, N, T, seed = 6, 7, 8, 6, 11
    args = _args(episode_length=T, n_rollout_threads=N)
    env = DcmlVecEnv(N, W, seed=seed)
    runner = DcmlRunner(dict(all_args=args, envs=env, num_agents=L, device=torch.device("cuda")))
    sd = mo.make_weights(5, L)
    runner.policy.transformer.load_state_dict(mo.full_state_dict(sd, L))
    runner.policy.transformer.rollout_seed = 777
    runner.policy.transformer.rollout_step = 0
    runner.warmup()
    for step in range(T):
        values, actions, logp, _, _ = runner.collect(step)
        obs, _, rew, dones, delay, pay, ava = env.step_device(actions)
        runner.insert((obs, env.sob, rew, dones, ava, values, actions, logp))
    runner.compute()
    b = runner.buffer
    # ---- the same rollout from the oracle's pieces -----------------------------------------------------
    ora = eo.DcmlEnvOracle(N, W, env.base_workload, seed=seed)
    o_obs, o_sob, o_ava = ora.reset()
    rewards, masks, vals = np.zeros((T, N, L, 1), np.float32), np.ones((T + 1, N, L, 1), np.float32), np.zeros((T + 1, N, L, 1), np.float32)
    for step in range(T):
        np.testing.assert_allclose(b.obs[step].cpu().numpy(), o_obs, rtol=1e-5, atol=1e-6)
        u = px.policy_uniforms(777, step, N, L)
        z = px.policy_normals(777, step, N, L - 1)
        vo, ao, lpo = mo.get_actions(sd, o_obs, o_ava, L, deterministic=False, u=u, z=z)
        a_gpu = b.actions[step].cpu().numpy().reshape(N, L)
        a_ora = ao.numpy().reshape(N, L)
        assert (a_gpu[:, :-1] == a_ora[:, :-1]).all(), "sampled discrete actions differ at step %d" % step
        np.testing.assert_allclose(a_gpu[:, -1], a_ora[:, -1], atol=1e-4)
        np.testing.assert_allclose(b.action_log_probs[step].cpu().numpy().reshape(N, L), lpo.numpy().reshape(N, L), atol=1e-4)
        np.testing.assert_allclose(b.value_preds[step].cpu().numpy().reshape(N, L), vo.numpy().reshape(N, L), atol=1e-4)
        vals[step] = vo.numpy().reshape(N, L, 1)
        # feed the oracle env the GPU path's actions so that a last-ulp difference in the Gaussian agent cannot fork the rollout
        o_obs, o_sob, o_rew, o_done, _, o_ava = ora.step(a_gpu.reshape(N, L, 1))
        np.testing.assert_allclose(b.rewards[step].cpu().numpy(), np.asarray(o_rew, np.float32).reshape(N, L, 1), rtol=1e-5)
        rewards[step] = np.asarray(o_rew, np.float32).reshape(N, L, 1)
        masks[step + 1] = 1.0 - np.asarray(o_done, np.float32).reshape(N, L, 1)
        np.testing.assert_array_equal(b.masks[step + 1].cpu().numpy(), masks[step + 1])
        np.testing.assert_array_equal(b.active_masks[step + 1].cpu().numpy(), np.ones((N, L, 1), np.float32))
    vals[T] = mo.get_values(sd, o_obs, L).numpy().reshape(N, L, 1)
    # compute(): GAE through the (freshly initialised) ValueNorm, base_runner.py:286-325 + shared_buffer.py:207-238
    adv_o, ret_o, _ = po.compute_returns(rewards, vals, masks, vals[T], vn=po.ValueNormState(), gamma=args.gamma, lam=args.gae_lambda)
    np.testing.assert_allclose(b.returns[:-1].cpu().numpy(), ret_o[:-1], rtol=2e-4, atol=2e-4)
    np.testing.assert_allclose(b.advantages.cpu().numpy(), adv_o, rtol=2e-4, atol=2e-4)
    # train(): finite losses, weights move, buffer rolled over (shared_buffer.py:188-198)
    w0 = runner.policy.transformer.flat_parameters().clone()
    last_obs = b.obs[-1].clone()
    info = runner.train()
    assert all(np.isfinite(v) for v in info.values()), info
    assert float((runner.policy.transformer.flat_parameters() - w0).abs().max()) > 0
    assert torch.equal(b.obs[0], last_obs)


def test_runner_run_loop(cuda_device):
    from mat_dcml_b200 import DcmlRunner, DcmlVecEnv
    W, L, N, T = 10, 11, 64, 10
    args = _args(episode_length=T, n_rollout_threads=N, num_env_steps=3 * T * N)
    runner = DcmlRunner(dict(all_args=args, envs=DcmlVecEnv(N, W, seed=2), num_agents=L, device=torch.device("cuda")))
    logs = runner.run()
    assert len(logs) == 3
    for i, info in enumerate(logs):
        assert info["total_num_steps"] == (i + 1) * T * N
        assert np.isfinite(info["value_loss"]) and np.isfinite(info["policy_loss"]) and np.isfinite(info["average_step_rewards"])
    # done episodes were seen (the simulator ends an episode with probability 1/2 per step ... SingleProcess.py:262-267)
    assert any("aver_episode_rewards" in info for info in logs)
    lrs = [g["lr"] for g in runner.policy.optimizer.param_groups]
    assert lrs[0] < args.lr          # linear decay applied (transformer_policy.py:108-114)


def test_runner_eval_matches_oracle_composition(cuda_device):
    """DCMLRunner.eval (dcml_runner.py:319-448): deterministic stride-2 decisions on preset contexts."""
    from mat_dcml_b200 import DcmlRunner, DcmlVecEnv
    W, L, N, steps, seed = 6, 7, 16, 12, 5
    args = _args(episode_length=4, n_rollout_threads=N)
    sd = mo.make_weights(9, L)
    R = np.linspace(5.0, 40.0, N)
    C = np.linspace(0.5, 0.05, N)

    def fresh():
        env = DcmlVecEnv(N, W, seed=seed)
        runner = DcmlRunner(dict(all_args=args, envs=env, num_agents=L, device=torch.device("cuda")))
        runner.policy.transformer.load_state_dict(mo.full_state_dict(sd, L))
        return env, runner

    # step-by-step: GPU decisions against the oracle's, oracle env driven by the GPU actions
    env, runner = fresh()
    ora = eo.DcmlEnvOracle(N, W, env.base_workload, seed=seed)
    obs, _, ava = env.reset_device()
    o_obs, _, o_ava = ora.reset()
    one = np.zeros((3, N)); fin = np.zeros(4)
    same_rows = 0
    for i in range(steps):
        act, _ = runner.policy.act(None, obs.reshape(N * L, -1), None, None, ava.reshape(N * L, -1), deterministic=True, stride=2)
        _, ao, lpo = mo.get_actions(sd, o_obs, o_ava, L, deterministic=True, stride=2)
        a_gpu, a_ora = act.cpu().numpy().reshape(N, L), ao.numpy().reshape(N, L)
        diff = a_gpu[:, :-1] != a_ora[:, :-1]
        same_rows += int((~diff.any(1)).sum())
        lpw = lpo.numpy().reshape(N, L)[:, :-1].astype(np.float64)
        with np.errstate(divide="ignore", invalid="ignore"):
            gap = np.where(lpw >= 0, np.inf, np.abs(lpw - np.log(-np.expm1(lpw))))      # the oracle's top-2 logit gap
        for e in np.nonzero(diff.any(1))[0]:
            assert gap[e, int(np.argmax(diff[e]))] < 1e-4, "deterministic action differs from the oracle away from a tie"
        obs, _, rew, dones, delay, pay, ava = env.step_device(act.reshape(N, L, 1))
        o_obs, _, o_rew, o_done, o_info, o_ava = ora.step(a_gpu.reshape(N, L, 1))
        np.testing.assert_allclose(rew.cpu().numpy(), np.asarray(o_rew)[:, 0, 0], rtol=1e-5)
        one[0] += np.asarray(o_rew)[:, 0, 0]
        one[1] += np.array([inf[0]["delay"] for inf in o_info]); one[2] += np.array([inf[0]["payment"] for inf in o_info])
        d = np.asarray(o_done)[:, 0].astype(bool)
        fin += [d.sum(), one[0][d].sum(), one[1][d].sum(), one[2][d].sum()]
        one[:, d] = 0.0
    assert same_rows >= 0.97 * steps * N
    assert fin[0] > 0
    # the runner's own loop on a fresh env of the same seed reports the same finished-episode means
    env2, runner2 = fresh()
    with pytest.raises(ValueError):
        runner2.eval(steps, stride=2, eval_envs=env2)          # never on the training envs
    r, dl, pay, t_dec = runner2.eval(steps, stride=2, eval_envs=DcmlVecEnv(N, W, seed=seed))
    np.testing.assert_allclose([r, dl, pay], fin[1:] / fin[0], rtol=1e-5)
    assert t_dec > 0
    # preset contexts (Env(preset=True) / modify_preset): the evaluation runs from the given R / C
    env3, runner3 = fresh()
    env3.reset_device()
    env3.set_context(R=R, C=C)
    o3 = eo.DcmlEnvOracle(N, W, env3.base_workload, seed=seed)
    o3.reset(); o3.set_context(R=R, C=C)
    obs3, _, ava3 = env3.obs, None, env3.ava
    np.testing.assert_allclose(obs3.cpu().numpy(), o3._observe_all()[0], rtol=1e-5, atol=1e-6)


def test_preset_benchmark_loop(cuda_device):
    """run_benchmark (DCML_MAT_ALT_Benchmark.py:115-140): every env gets its preset row, the decisions are the
    oracle's deterministic stride-10 decisions, and the sweep means are reproducible."""
    from conftest import golden
    from mat_dcml_b200 import DcmlVecEnv, TransformerPolicy
    from mat_dcml_b200.benchmark import PresetSamples, run_benchmark, sweep
    g = golden("preset_sample1")
    samples = PresetSamples(g["master"], g["worker_prs"], g["disable_rates"])
    W, L, N = 100, 101, 8
    args = _args(episode_length=4, n_rollout_threads=N)
    sd = mo.make_weights(11, L)

    def fresh(seed=6):
        env = DcmlVecEnv(N, W, seed=seed)
        pol = TransformerPolicy(args, [7], [W + 2], env.action_space[0], L, device=torch.device("cuda"))
        pol.transformer.load_state_dict(mo.full_state_dict(sd, L))
        return env, pol

    env, pol = fresh()
    env.reset_device()
    obs, _, ava = samples.apply(env, start=4)
    ora = eo.DcmlEnvOracle(N, W, env.base_workload, seed=6)
    ora.reset()
    idx = 4 + np.arange(N)
    o_obs, _, o_ava = ora.set_context(R=samples.master_status[idx, 0], C=samples.master_status[idx, 1],
                                      Pr=samples.worker_Prs[idx], k_dis=samples.disable_rates[idx])
    # the oracle keeps its disabled SET when only the count changes; compare what does not depend on the set
    ob = obs.cpu().numpy()
    np.testing.assert_allclose(ob[:, :, :2], o_obs[:, :, :2], rtol=1e-5, atol=1e-6)          # normalised R, C on every row
    assert ((ava.cpu().numpy()[:, :W, 1] == 0).sum(1) == samples.disable_rates[idx]).all()   # preset disable counts
    env, pol = fresh()                           # same seed + same call sequence = same named draws (a reset never rewinds)
    r1 = run_benchmark(pol, env, samples, n_episodes=20, stride=10)
    assert r1["rewards"].numel() == 20 and np.isfinite([r1["reward"], r1["ct"], r1["payment"]]).all() and r1["decision_s"] > 0
    ratio = (r1["rewards"] / (-99.0 * r1["cts"] - r1["payments"])).cpu().numpy()       # DCML_ENV_Functions.py:6-17; x1.5 when N = 0
    assert (np.isclose(ratio, 1.0, rtol=1e-5) | np.isclose(ratio, 1.5, rtol=1e-5)).all()
    env2, pol2 = fresh()
    r2 = run_benchmark(pol2, env2, samples, n_episodes=20, stride=10)
    np.testing.assert_array_equal(r1["rewards"].cpu().numpy(), r2["rewards"].cpu().numpy())
    cts, pays, rews = sweep(pol, env, samples, [{"disable_rate": 8}, {"disable_rate": 64}], n_episodes=16)
    assert cts.shape == (2, 1) and pays.shape == (2, 1) and np.isfinite(rews).all()
    assert (samples.disable_rates == g["disable_rates"]).all()       # the sweep works on copies


def test_buffer_generator_vs_reference(cuda_device, monkeypatch):
    """SharedReplayBuffer.insert / compute_returns / feed_forward_generator_transformer / after_update against the reference's
    buffer on the same content with the same injected permutation (fixture `buffer`, shared_buffer.py:118-151,188-198,
    240-314): the 12-tuple of every minibatch -- row-to-sample grouping, agent order, share_obs replication -- bit for bit.
    use_gae=False is accepted and ignored like in the reference (compute_returns only has GAE branches)."""
    from conftest import golden
    from mat_dcml_b200 import SharedReplayBuffer, Action_Space
    g = golden("buffer")
    T, N, L, W = int(g["T"]), int(g["N"]), int(g["L"]), int(g["W"])
    buf = SharedReplayBuffer(_args(episode_length=T, n_rollout_threads=N, use_valuenorm=False, use_popart=False, use_gae=False),
                             L, [7], [W + 2], Action_Space(2, semi_index=-1, extra=True), "DCML", device=cuda_device)
    dev = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(cuda_device)
    buf._share_obs[0].copy_(dev(g["first_share_obs"][:, 0])); buf.obs[0].copy_(dev(g["first_obs"]))
    buf.available_actions[0].copy_(dev(g["first_ava"]))
    for t in range(T):
        s = lambda k: g["s%d_%s" % (t, k)]
        # host numpy inputs in the reference's replicated layout (share_obs (N, L, D), rewards / masks (N, L, 1))
        buf.insert(s("share_obs"), s("obs"), None, None, s("actions"), s("logp"), s("values"), s("rewards"), s("masks"), None, None,
                   s("ava"))
    buf.compute_returns(dev(g["next_value"]), None)
    perm = torch.from_numpy(g["perm"]).to(cuda_device)
    monkeypatch.setattr(torch, "randperm", lambda n, **k: perm.clone())
    names = ("share_obs", "obs", "rnn", "rnn_c", "actions", "value_preds", "returns", "masks", "active", "old_logp", "adv", "ava")
    n_mb = 0
    for i, sample in enumerate(buf.feed_forward_generator_transformer(dev(g["adv"]), num_mini_batch=3)):
        n_mb += 1
        assert len(sample) == 12 and sample[2] is None and sample[3] is None
        for k, v in zip(names, sample):
            if k.startswith("rnn"):
                continue
            want = g["mb%d_%s" % (i, k)]
            assert tuple(v.shape) == want.shape, (k, v.shape, want.shape)
            np.testing.assert_array_equal(v.cpu().numpy(), want, err_msg="minibatch %d %s" % (i, k))
    assert n_mb == 3
    monkeypatch.undo()
    buf.after_update()
    np.testing.assert_array_equal(buf.obs[0].cpu().numpy(), g["after_obs0"])
    np.testing.assert_array_equal(buf._share_obs[0].cpu().numpy(), g["after_share0"])
    np.testing.assert_array_equal(buf.masks[0].cpu().numpy(), g["after_masks0"])
    np.testing.assert_array_equal(buf.available_actions[0].cpu().numpy(), g["after_ava0"])
