"""GPU parity of the DCML simulator and the GAE kernel, through the C ABI, against reference
outputs (golden fixtures) and the CPU oracle."""
import numpy as np
import pytest
import torch

from conftest import golden
from oracle import env_oracle as eo
from oracle import ppo_oracle as po

pytestmark = pytest.mark.gpu
RTOL = 1e-5          # BASELINE.json: env outputs and rewards within 1e-5 relative in fp32


def _close(got, want, name=""):
    np.testing.assert_allclose(np.asarray(got, np.float64), np.asarray(want, np.float64), rtol=RTOL, atol=1e-6, err_msg=name)


@pytest.mark.parametrize("case", ["env_W20", "env_W100", "env_W50_wide", "env_W100_wide"])
def test_env_vs_reference(case, cuda_device):
    from mat_dcml_b200 import DcmlVecEnv
    g = golden(case)
    W, B, steps = int(g["W"]), int(g["B"]), int(g["steps"])
    env = DcmlVecEnv(B, W, g["base"], seed=int(g["seed"]), double_reset=True)
    obs, sob, ava = env.reset()
    assert obs.shape == (B, W + 1, 7) and sob.shape == (B, W + 1, W + 2) and ava.shape == (B, W + 1, 2)
    assert obs.dtype == np.float64 and ava.dtype == np.int64
    _close(obs, g["obs0"], "obs0"); _close(sob[:, 0], g["sob0"], "sob0")
    np.testing.assert_array_equal(ava, g["ava0"])
    for t in range(1, steps + 1):
        obs, sob, rew, dones, infos, ava = env.step(g["act%d" % t])
        assert rew.shape == (B, W + 1, 1) and dones.shape == (B, W + 1) and dones.dtype == bool
        _close([i[0]["delay"] for i in infos], g["delay%d" % t], "delay")
        _close([i[0]["payment"] for i in infos], g["pay%d" % t], "payment")
        _close(rew[:, 0, 0], g["rew%d" % t], "reward")
        np.testing.assert_array_equal(dones[:, 0], g["done%d" % t])
        _close(obs, g["obs%d" % t], "obs"); _close(sob[:, 0], g["sob%d" % t], "sob")
        np.testing.assert_array_equal(ava, g["ava%d" % t])


@pytest.mark.parametrize("W,B,dbl", [(10, 37, True), (50, 9, False), (200, 3, True)])
def test_env_vs_oracle_random(W, B, dbl, cuda_device):
    from mat_dcml_b200 import DcmlVecEnv
    base = eo.synthetic_workloads(W, seed=W)
    env = DcmlVecEnv(B, W, base, seed=77, double_reset=dbl, env_offset=1000)
    ora = eo.DcmlEnvOracle(B, W, base, seed=77, double_reset=dbl, env0=1000)
    o1, s1, a1 = env.reset()
    o2, s2, a2 = ora.reset()
    _close(o1, o2); _close(s1, s2); np.testing.assert_array_equal(a1, a2)
    rng = np.random.RandomState(W)
    for t in range(4):
        sel = (rng.uniform(size=(B, W)) < 0.5).astype(np.float32) * a2[:, :W, 1]
        if t == 1:
            sel[0] = 0
        act = np.concatenate([sel, rng.uniform(-0.2, 1.3, size=(B, 1)).astype(np.float32)], 1)[:, :, None]
        o1, s1, r1, d1, i1, a1 = env.step(act)
        o2, s2, r2, d2, i2, a2 = ora.step(act)
        _close(r1, r2, "reward"); np.testing.assert_array_equal(d1, d2)
        _close([i[0]["delay"] for i in i1], [i[0]["delay"] for i in i2])
        _close([i[0]["payment"] for i in i1], [i[0]["payment"] for i in i2])
        _close(o1, o2, "obs"); _close(s1, s2, "sob"); np.testing.assert_array_equal(a1, a2)


def test_env_shannon_vs_reference_and_oracle(cuda_device):
    """Shannon link rates: reset(shannon_enable=True) + the following step of the reference."""
    from mat_dcml_b200 import DcmlVecEnv
    g = golden("env_W12_shannon")
    W, B = int(g["W"]), int(g["B"])
    env = DcmlVecEnv(B, W, g["base"], seed=int(g["seed"]), shannon=True)
    obs, sob, ava = env.reset()
    assert sob.shape == (B, W + 1, 2 * W + 2)
    _close(obs, g["obs0"], "obs0"); _close(sob[:, 0], g["sob0"], "sob0")
    np.testing.assert_array_equal(ava, g["ava0"])
    _, _, rew, dones, infos, _ = env.step(g["act1"])
    _close([i[0]["delay"] for i in infos], g["delay1"]); _close([i[0]["payment"] for i in infos], g["pay1"])
    _close(rew[:, 0, 0], g["rew1"]); np.testing.assert_array_equal(dones[:, 0], g["done1"])
    # several steps against the oracle with Shannon rates on every reset
    W, B = 30, 11
    env = DcmlVecEnv(B, W, seed=9, shannon=True)
    ora = eo.DcmlEnvOracle(B, W, env.base_workload, seed=9, shannon=True)
    o1, s1, a1 = env.reset(); o2, s2, a2 = ora.reset()
    _close(o1, o2); _close(s1, s2); np.testing.assert_array_equal(a1, a2)
    rng = np.random.RandomState(1)
    for t in range(3):
        act = np.concatenate([(rng.uniform(size=(B, W)) < 0.5).astype(np.float32) * a2[:, :W, 1],
                              rng.uniform(0.1, 1.0, size=(B, 1)).astype(np.float32)], 1)[:, :, None]
        o1, s1, r1, d1, i1, a1 = env.step(act); o2, s2, r2, d2, i2, a2 = ora.step(act)
        _close(r1, r2); _close(o1, o2); _close(s1, s2); np.testing.assert_array_equal(d1, d2)


def test_env_single_agent_binary_vs_reference(cuda_device):
    """DcmlVecEnv(multi_agent=False).reset(binary=True) + steps against the reference's Env(multi_agent=False)
    (…SingleProcess.py:42-47,164-172,272-274; DCML_Basic_Env.py:14-17)."""
    from mat_dcml_b200 import DcmlVecEnv
    g = golden("env_W16_single_binary")
    W, B, steps = int(g["W"]), int(g["B"]), int(g["steps"])
    L = W + 1
    env = DcmlVecEnv(B, W, g["base"], seed=int(g["seed"]), double_reset=False, multi_agent=False)
    assert env.observation_space == [[7 * L]] and env.action_space[0].mixed and env.action_space[0].high == W
    obs, sob, ava = env.reset(binary=True)
    assert obs.shape == (B, 7 * L) and sob.shape == (B, 64 + W) and ava.shape == (B, L, 2)
    _close(obs, g["obs0"], "obs0"); _close(sob, g["sob0"], "sob0")
    np.testing.assert_array_equal(ava, g["ava"][:, 0])
    for t in range(steps):
        obs, sob, rew, dones, infos, ava = env.step(g["act"][:, t])
        assert obs.shape == (B, 7 * L) and sob.shape == (B, W + 2) and rew.shape == (B, 1) and dones.shape == (B, 1)
        _close(rew[:, 0], g["rew"][:, t], "reward")
        _close([i[0]["delay"] for i in infos], g["delay"][:, t]); _close([i[0]["payment"] for i in infos], g["pay"][:, t])
        np.testing.assert_array_equal(dones[:, 0], g["done"][:, t])
        _close(obs, g["obs"][:, t], "obs"); _close(sob, g["sob"][:, t], "sob")
        np.testing.assert_array_equal(ava, g["ava"][:, t + 1])
    # the multi-agent contract after a binary reset: the wider row replicated over the agents
    env2 = DcmlVecEnv(B, W, g["base"], seed=int(g["seed"]), double_reset=False)
    o2, s2, a2 = env2.reset(binary=True)
    assert s2.shape == (B, L, 64 + W)
    _close(o2.reshape(B, 7 * L), g["obs0"]); _close(s2[:, 3], g["sob0"])


def test_env_repeated_reset_draws_a_new_context(cuda_device):
    """A second reset() is the reference's env.reset() on a live env: fresh context, workload tables kept, no replay."""
    from mat_dcml_b200 import DcmlVecEnv
    W, B = 12, 9
    env = DcmlVecEnv(B, W, seed=21)
    ora = eo.DcmlEnvOracle(B, W, env.base_workload, seed=21)
    o1, _, a1 = env.reset(); p1, _, b1 = ora.reset()
    _close(o1, p1); np.testing.assert_array_equal(a1, b1)
    act = np.concatenate([a1[:, :W, 1].astype(np.float32), np.full((B, 1), 0.6, np.float32)], 1)[:, :, None]
    env.step(act); ora.step(act)
    o2, _, a2 = env.reset(); p2, _, b2 = ora.reset()
    _close(o2, p2); np.testing.assert_array_equal(a2, b2)
    assert not np.allclose(o1[:, 0, :2], o2[:, 0, :2])            # new (R, C)
    act = np.concatenate([a2[:, :W, 1].astype(np.float32), np.full((B, 1), 0.6, np.float32)], 1)[:, :, None]
    _, _, r1, d1, _, _ = env.step(act); _, _, r2, d2, _, _ = ora.step(act)
    _close(r1, r2); np.testing.assert_array_equal(d1, d2)
    with pytest.raises(ValueError):
        env.set_context(n_disabled=np.full(B, W))


def test_env_preset_context(cuda_device):
    from mat_dcml_b200 import DcmlVecEnv
    W, B = 20, 5
    env = DcmlVecEnv(B, W, seed=3)
    env.reset()
    pr = np.full((B, W), 0.25)
    obs, sob, ava = env.set_context(R=2 ** 19, C=921.6, worker_pr=pr, n_disabled=np.array([0, 1, 8, 16, 16]))
    obs, ava = obs.cpu().numpy(), ava.cpu().numpy()
    np.testing.assert_allclose(obs[:, :, 0], (2 ** 19 - 1024) / (2 ** 20 - 1024), rtol=1e-6)
    np.testing.assert_allclose(obs[:, :, 1], (921.6 - 32) / (1024 - 32), rtol=1e-6)
    np.testing.assert_array_equal((ava[:, :W, 1] == 0).sum(1), [0, 1, 8, 16, 16])
    ora = eo.DcmlEnvOracle(B, W, env.base_workload, seed=3)
    ora.reset()
    ora.set_context(R=2 ** 19, C=921.6, Pr=pr, k_dis=np.array([0, 1, 8, 16, 16]), disabled=(ava[:, :W, 1] == 0))
    act = np.concatenate([ava[:, :W, 1], np.full((B, 1), 0.7, np.float32)], 1).astype(np.float32)
    _, _, r1, _, i1, _ = env.step(act)
    _, _, r2, _, i2, _ = ora.step(act)
    _close(r1, r2); _close([i[0]["delay"] for i in i1], [i[0]["delay"] for i in i2])


def test_env_full_size_properties(cuda_device):
    """C3 shape (16384 x 50): invariants of the simulator that hold at any size, and the oracle itself on three runs of
    envs taken from the start, the middle and the end of the batch (envs are independent: named draws by env index)."""
    from mat_dcml_b200 import DcmlVecEnv
    W, B = 50, 16384
    env = DcmlVecEnv(B, W, seed=5)
    starts = (0, B // 2 - 11, B - 21)
    oras = [eo.DcmlEnvOracle(21, W, env.base_workload, seed=5, env0=s0) for s0 in starts]
    obs, sob, ava = env.reset_device()
    for s0, ora in zip(starts, oras):
        o2, s2, a2 = ora.reset()
        _close(obs[s0:s0 + 21].cpu().numpy(), o2, "obs0"); _close(env.sob[s0:s0 + 21].cpu().numpy(), s2[:, 0], "sob0")
        np.testing.assert_array_equal(ava[s0:s0 + 21].cpu().numpy(), a2)
    dones = []
    for t in range(3):
        sel = ava[:, :W, 1] * (torch.rand(B, W, device="cuda") < 0.6)
        act = torch.cat([sel, torch.full((B, 1), 0.7, device="cuda")], 1)
        if t == 1:
            act[0, :W] = 0                                  # the N == 0 branch in the first tile
        obs, sob, rew, done, delay, pay, ava = env.step_device(act)
        for s0, ora in zip(starts, oras):
            o2, s2, r2, d2, i2, a2 = ora.step(act[s0:s0 + 21].cpu().numpy()[:, :, None])
            _close(rew[s0:s0 + 21].cpu().numpy(), r2[:, 0, 0], "reward")
            _close(delay[s0:s0 + 21].cpu().numpy(), [i[0]["delay"] for i in i2], "delay")
            _close(pay[s0:s0 + 21].cpu().numpy(), [i[0]["payment"] for i in i2], "payment")
            np.testing.assert_array_equal(done[s0:s0 + 21].cpu().numpy().astype(bool), d2[:, 0])
            _close(obs[s0:s0 + 21].cpu().numpy(), o2, "obs"); _close(env.sob[s0:s0 + 21].cpu().numpy(), s2[:, 0], "sob")
            np.testing.assert_array_equal(ava[s0:s0 + 21].cpu().numpy(), a2)
        n_sel = act[:, :W].sum(1)
        scale = torch.where(n_sel == 0, 1.5, 1.0)
        assert torch.allclose(rew, scale * (-99 * delay - pay), rtol=1e-5, atol=1e-4)
        assert bool((delay > 0).all()) and bool((pay >= 0).all()) and torch.isfinite(rew).all()
        k = (ava[:, :W, 1] == 0).sum(1)
        assert int(k.min()) >= 1 and int(k.max()) <= 40
        assert bool((obs[:, :W, 2:6][ava[:, :W, 1] == 0] == 1).all())
        assert bool((obs[:, W, 6] == 1.1).all()) and bool((ava[:, W] == 1).all())
        assert float(obs[:, :W, 6].max()) < 1.0 and float(obs.min()) >= 0
        dones.append(done.float().mean().item())
    assert abs(np.mean(dones) - 0.8) < 0.02


def _vn(stats, device):
    from mat_dcml_b200 import ValueNorm
    vn = ValueNorm(1, device=device)
    vn.stats.copy_(torch.as_tensor(stats))
    return vn


class _IeeeVN(po.ValueNormState):
    """ValueNorm with a correctly rounded sqrt.  torch-CPU's vectorised sqrt (what the reference
    runs) is off by one ulp on near-ties -- e.g. sqrt(1582.46484375f), the variance in the golden
    fixture -- whereas the kernel's sqrtf is IEEE; see DESIGN.md 'GAE'."""

    def denormalize(self, x):
        mean, var = self.mean_var()
        sd = np.sqrt(np.float32(var.item()))
        return (np.asarray(x, np.float32) * sd + np.float32(mean.item())).astype(np.float32)


def test_gae_vs_reference(cuda_device):
    """raw GAE: bit-exact with the reference; with ValueNorm: bit-exact with the IEEE-sqrt
    restatement and within 1e-5 relative of the reference (one-ulp sqrt difference, see _IeeeVN)."""
    from mat_dcml_b200 import SharedReplayBuffer, Action_Space
    import argparse
    g = golden("gae")
    for case in ("vn", "raw"):
        T, N, L, _ = g[case + "_rewards"].shape
        args = argparse.Namespace(episode_length=T, n_rollout_threads=N, hidden_size=64, recurrent_N=1, gamma=0.99,
                                  gae_lambda=0.95, use_gae=True, use_popart=False, use_valuenorm=case == "vn",
                                  use_proper_time_limits=False, algorithm_name="mat")
        buf = SharedReplayBuffer(args, L, [7], [L + 1], Action_Space(2, semi_index=-1), "DCML")
        buf.rewards.copy_(torch.from_numpy(g[case + "_rewards"]))
        buf.value_preds.copy_(torch.from_numpy(g[case + "_values_in"]))
        buf.masks.copy_(torch.from_numpy(g[case + "_masks"]))
        buf.compute_returns(g[case + "_next"], _vn(g["vn_stats"], "cuda") if case == "vn" else None)
        adv, ret = buf.advantages.cpu().numpy(), buf.returns.cpu().numpy()[:-1]
        if case == "raw":
            np.testing.assert_array_equal(adv, g[case + "_adv"])
            np.testing.assert_array_equal(ret, g[case + "_ret"][:-1])
        else:
            scale = np.abs(g[case + "_rewards"]).max()
            np.testing.assert_allclose(adv, g[case + "_adv"], rtol=1e-5, atol=1e-5 * scale)
            np.testing.assert_allclose(ret, g[case + "_ret"][:-1], rtol=1e-5, atol=1e-5 * scale)
            vn = _IeeeVN()
            vn.running_mean[:] = float(g["vn_stats"][0]); vn.running_mean_sq[:] = float(g["vn_stats"][1])
            vn.debiasing_term.fill_(float(g["vn_stats"][2]))
            a2, r2, _ = po.compute_returns(g[case + "_rewards"], g[case + "_values_in"], g[case + "_masks"], g[case + "_next"], vn)
            np.testing.assert_array_equal(adv, a2)
            np.testing.assert_array_equal(ret, r2[:-1])


@pytest.mark.parametrize("T,n", [(50, 8192 * 21), (7, 13), (3, 4097)])
def test_gae_vs_oracle(T, n, cuda_device):
    import ctypes
    from mat_dcml_b200 import _cabi
    rng = np.random.RandomState(n % 1000)
    r = (rng.standard_normal((T, n)) * 50 - 200).astype(np.float32)
    v = rng.standard_normal((T + 1, n)).astype(np.float32)
    m = (rng.uniform(size=(T + 1, n)) > 0.8).astype(np.float32)
    vn = _IeeeVN()
    vn.update(torch.from_numpy((rng.standard_normal((64, 1)) * 40 - 150).astype(np.float32)))
    adv, ret, _ = po.compute_returns(r[..., None, None], v[..., None, None], m[..., None, None], v[-1][..., None, None], vn)
    d = lambda x: torch.from_numpy(x).cuda()
    dr, dv, dm, ds = d(r), d(v), d(m), d(vn.stats())
    da, dret = torch.zeros_like(dr), torch.zeros_like(dv)
    _cabi.check(_cabi.lib().matas_gae(_cabi.ptr(dr), _cabi.ptr(dv), _cabi.ptr(dm), _cabi.ptr(ds), 0.99, 0.95, T, n,
                                      _cabi.ptr(da), _cabi.ptr(dret), _cabi.stream_ptr()))
    np.testing.assert_array_equal(da.cpu().numpy(), adv[..., 0, 0])
    np.testing.assert_array_equal(dret.cpu().numpy()[:-1], ret[:-1, :, 0, 0])


@pytest.mark.parametrize("tag,kw", [("fx", dict(fixed=True)), ("sa", dict(standalone=True))])
def test_env_variants_vs_reference_and_oracle(tag, kw, cuda_device):
    """Env(fixed=True) and step(standalone=True) of the reference (fixture), then a larger ragged batch against the oracle."""
    from mat_dcml_b200 import DcmlVecEnv
    g = golden("env_W16_variants")
    W, B, steps = int(g["W"]), int(g["B"]), int(g["steps"])
    env = DcmlVecEnv(B, W, g["base"], seed=int(g["seed"]), double_reset=False, **kw)
    obs, sob, ava = env.reset()
    _close(obs, g[tag + "_obs"][:, 0], "obs0"); np.testing.assert_array_equal(ava, g[tag + "_ava"][:, 0])
    for t in range(1, steps + 1):
        obs, sob, rew, dones, infos, ava = env.step(g[tag + "_act"][:, t - 1])
        _close([i[0]["delay"] for i in infos], g[tag + "_delay"][:, t - 1], "delay")
        _close([i[0]["payment"] for i in infos], g[tag + "_pay"][:, t - 1], "payment")
        _close(rew[:, 0, 0], g[tag + "_rew"][:, t - 1], "reward")
        np.testing.assert_array_equal(dones[:, 0], g[tag + "_done"][:, t - 1])
        _close(obs, g[tag + "_obs"][:, t], "obs"); _close(sob[:, 0], g[tag + "_sob"][:, t], "sob")
        np.testing.assert_array_equal(ava, g[tag + "_ava"][:, t])
    W, B = 37, 13
    env = DcmlVecEnv(B, W, seed=4, **kw)
    ora = eo.DcmlEnvOracle(B, W, env.base_workload, seed=4, **kw)
    o1, s1, a1 = env.reset(); o2, s2, a2 = ora.reset()
    rng = np.random.RandomState(2)
    for t in range(3):
        act = np.concatenate([(rng.uniform(size=(B, W)) < 0.5).astype(np.float32) * a2[:, :W, 1],
                              rng.uniform(0.1, 1.0, size=(B, 1)).astype(np.float32)], 1)[:, :, None]
        o1, s1, r1, d1, i1, a1 = env.step(act); o2, s2, r2, d2, i2, a2 = ora.step(act)
        _close(r1, r2); _close(o1, o2); _close(s1, s2); np.testing.assert_array_equal(d1, d2)


def test_philox_device_known_answers(cuda_device):
    """csrc/common.cuh draws the same Philox4x32-10 words as Random123's known-answer vectors and as oracle/philox.py."""
    from mat_dcml_b200 import _cabi
    from oracle import philox as px
    from test_oracle_golden import PHILOX_KAT
    rows = [list(c) + list(k) for c, k, _ in PHILOX_KAT]
    rng = np.random.RandomState(0)
    extra = rng.randint(0, 2 ** 32, size=(1000, 6), dtype=np.uint64)
    inp32 = torch.from_numpy(np.concatenate([np.array(rows, np.uint64), extra]).astype(np.uint32).view(np.int32)).cuda()
    out = torch.empty(inp32.shape[0], 4, dtype=torch.int32, device="cuda")
    _cabi.check(_cabi.lib().matas_philox4x32_10(_cabi.ptr(inp32), _cabi.ptr(out), inp32.shape[0], _cabi.stream_ptr()))
    got = out.cpu().numpy().view(np.uint32)
    for i, (_, _, want) in enumerate(PHILOX_KAT):
        assert tuple(int(x) for x in got[i]) == want
    e = extra.astype(np.uint64)
    for i in range(0, 1000, 37):
        w = px.philox4x32(int(e[i, 0]), int(e[i, 1]), int(e[i, 2]), int(e[i, 3]), seed=int(e[i, 4]) | (int(e[i, 5]) << 32))
        assert [int(x) for x in got[3 + i]] == [int(x) for x in w]
