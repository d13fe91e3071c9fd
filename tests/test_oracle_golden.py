"""Pin the CPU oracle (oracle/) against outputs of the unmodified reference
(fixtures written by tests/golden/make_golden.py).  CPU only."""
import numpy as np
import pytest
import torch

from conftest import golden
from oracle import env_oracle as eo
from oracle import mat_oracle as mo
from oracle import ppo_oracle as po
from oracle import philox as px

POLICY_CASES = ["policy_L21_trained", "policy_L21_init", "policy_L101_trained"]


def _weights(g, semi=True):
    return mo.make_weights(int(g["wseed"]), int(g["L"]), profile=str(g["profile"]), semi=semi)


@pytest.mark.parametrize("case", POLICY_CASES)
def test_policy_deterministic_strides(case):
    g = golden(case)
    L = int(g["L"])
    sd = _weights(g)
    np.testing.assert_allclose(mo.get_values(sd, g["obs"], L).numpy(), g["values"], atol=2e-6, rtol=1e-5)
    for s in (1, 2, 10, L):
        v, a, lp = mo.get_actions(sd, g["obs"], g["ava"], L, deterministic=True, stride=s)
        np.testing.assert_array_equal(a.numpy()[: -1], g["det%d_actions" % s][: -1])
        np.testing.assert_allclose(a.numpy(), g["det%d_actions" % s], atol=2e-6, rtol=1e-5)
        np.testing.assert_allclose(lp.numpy(), g["det%d_logp" % s], atol=2e-6, rtol=1e-5)
        np.testing.assert_allclose(v.numpy(), g["det%d_values" % s], atol=2e-6, rtol=1e-5)


@pytest.mark.parametrize("case", POLICY_CASES)
def test_policy_evaluate(case):
    g = golden(case)
    L = int(g["L"])
    sd = _weights(g)
    for tag in ("det", "smp"):
        acts = g["det1_actions"] if tag == "det" else g["smp_actions"]
        obs = torch.from_numpy(g["obs"]).float().reshape(-1, L, 7)
        _, rep = mo.encoder(sd, obs)
        lp, ent = mo.semi_discrete_parallel_act(sd, rep, torch.from_numpy(acts).float().reshape(-1, L, 1),
                                                torch.from_numpy(g["ava"]).float().reshape(-1, L, 2))
        np.testing.assert_allclose(lp.reshape(-1, 1).numpy(), g["ev_%s_logp" % tag], atol=2e-6, rtol=1e-5)
        np.testing.assert_allclose(ent.reshape(-1, 1).numpy(), g["ev_%s_entropy" % tag], atol=2e-6, rtol=1e-5)
        v, lp2, es = mo.evaluate_actions(sd, g["obs"], acts, g["ava"], L, g["active"])
        np.testing.assert_allclose(v.numpy(), g["ev_%s_values" % tag], atol=2e-6, rtol=1e-5)
        np.testing.assert_allclose(es.numpy(), g["ev_%s_entropy_scalar" % tag], atol=2e-6, rtol=1e-5)
    # the reference's sampled log-probs equal its teacher-forced ones
    np.testing.assert_allclose(g["smp_logp"], g["ev_smp_logp"], atol=2e-6, rtol=1e-5)


def test_policy_discrete_variant():
    g = golden("policy_L13_discrete")
    L = int(g["L"])
    sd = _weights(g, semi=False)
    v, a, lp = mo.get_actions(sd, g["obs"], g["ava"], L, deterministic=True, mode="discrete")
    np.testing.assert_array_equal(a.numpy(), g["det4_actions"])
    np.testing.assert_allclose(lp.numpy(), g["det4_logp"], atol=2e-6, rtol=1e-5)
    for tag in ("det", "smp"):
        acts = g["det4_actions"] if tag == "det" else g["smp_actions"]
        v, lp2, es = mo.evaluate_actions(sd, g["obs"], acts, g["ava"], L, g["active"], mode="discrete")
        np.testing.assert_allclose(lp2.numpy(), g["ev_%s_logp" % tag], atol=2e-6, rtol=1e-5)
        np.testing.assert_allclose(es.numpy(), g["ev_%s_entropy_scalar" % tag], atol=2e-6, rtol=1e-5)


def test_policy_stochastic_injection_is_self_consistent():
    """The uniform-injection decode (the only part with no reference counterpart) must score
    its own samples exactly like teacher forcing does, and follow the injected uniforms."""
    g = golden("policy_L21_trained")
    L, B = int(g["L"]), int(g["B"])
    sd = _weights(g)
    from oracle import philox as px
    u = px.policy_uniforms(5, 3, B, L)
    z = px.policy_normals(5, 3, B, L - 1)
    v, a, lp = mo.get_actions(sd, g["obs"], g["ava"], L, deterministic=False, u=u, z=z)
    _, lp2, _ = mo.evaluate_actions(sd, g["obs"], a.numpy(), g["ava"], L)
    np.testing.assert_allclose(lp.numpy(), lp2.numpy(), atol=3e-6, rtol=1e-5)
    a = a.numpy().reshape(B, L)
    ava = g["ava"].reshape(B, L, 2)
    assert np.all(a[:, :-1][ava[:, :-1, 1] == 0] == 0)
    assert 0.05 < a[:, :-1][ava[:, :-1, 1] == 1].mean() < 0.95


@pytest.mark.parametrize("case", ["env_W20", "env_W100", "env_W50_wide", "env_W100_wide"])
def test_env_matches_reference(case):
    g = golden(case)
    W, B, steps = int(g["W"]), int(g["B"]), int(g["steps"])
    env = eo.DcmlEnvOracle(B, W, g["base"], seed=int(g["seed"]), double_reset=True)
    obs, sob, ava = env.reset()
    np.testing.assert_allclose(obs, g["obs0"], rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(sob[:, 0], g["sob0"], rtol=1e-12, atol=1e-14)
    np.testing.assert_array_equal(ava, g["ava0"])
    for t in range(1, steps + 1):
        obs, sob, rew, dones, infos, ava = env.step(g["act%d" % t])
        np.testing.assert_allclose([i[0]["delay"] for i in infos], g["delay%d" % t], rtol=1e-12)
        np.testing.assert_allclose([i[0]["payment"] for i in infos], g["pay%d" % t], rtol=1e-12)
        np.testing.assert_allclose(rew[:, 0, 0], g["rew%d" % t], rtol=1e-12)
        np.testing.assert_array_equal(dones[:, 0], g["done%d" % t])
        np.testing.assert_allclose(obs, g["obs%d" % t], rtol=1e-12, atol=1e-14)
        np.testing.assert_allclose(sob[:, 0], g["sob%d" % t], rtol=1e-12, atol=1e-14)
        np.testing.assert_array_equal(ava, g["ava%d" % t])


def test_env_single_agent_binary_matches_reference():
    """Env(multi_agent=False) + reset(binary=True): the reference's flat single-learner contract is a reshape of the same
    simulator; the binary reset changes share_obs (64 + W wide) and observation features 0 / 1 (fixture env_W16_single_binary)."""
    g = golden("env_W16_single_binary")
    W, B, steps = int(g["W"]), int(g["B"]), int(g["steps"])
    L = W + 1
    env = eo.DcmlEnvOracle(B, W, g["base"], seed=int(g["seed"]), double_reset=False)
    obs, sob, ava = env.reset(binary=True)
    np.testing.assert_allclose(obs.reshape(B, L * 7), g["obs0"], rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(sob[:, 0], g["sob0"], rtol=1e-12, atol=1e-14)
    np.testing.assert_array_equal(ava, g["ava"][:, 0])
    for t in range(steps):
        act = g["act"][:, t]
        obs, sob, rew, dones, infos, ava = env.step(act[:, :, None])
        np.testing.assert_allclose(rew[:, 0, 0], g["rew"][:, t], rtol=1e-12)
        np.testing.assert_allclose([i[0]["delay"] for i in infos], g["delay"][:, t], rtol=1e-12)
        np.testing.assert_array_equal(dones[:, 0], g["done"][:, t])
        np.testing.assert_allclose(obs.reshape(B, L * 7), g["obs"][:, t], rtol=1e-12, atol=1e-14)
        np.testing.assert_allclose(sob[:, 0], g["sob"][:, t], rtol=1e-12, atol=1e-14)
        np.testing.assert_array_equal(ava, g["ava"][:, t + 1])


def test_env_shannon_matches_reference():
    g = golden("env_W12_shannon")
    W, B = int(g["W"]), int(g["B"])
    env = eo.DcmlEnvOracle(B, W, g["base"], seed=int(g["seed"]), shannon=True)
    obs, sob, ava = env.reset()
    np.testing.assert_allclose(obs, g["obs0"], rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(sob[:, 0], g["sob0"], rtol=1e-12)
    np.testing.assert_array_equal(ava, g["ava0"])
    _, _, rew, dones, infos, _ = env.step(g["act1"])
    np.testing.assert_allclose([i[0]["delay"] for i in infos], g["delay1"], rtol=1e-12)
    np.testing.assert_allclose([i[0]["payment"] for i in infos], g["pay1"], rtol=1e-12)
    np.testing.assert_allclose(rew[:, 0, 0], g["rew1"], rtol=1e-12)
    np.testing.assert_array_equal(dones[:, 0], g["done1"])


def _vn_from(stats):
    vn = po.ValueNormState()
    vn.running_mean[:] = float(stats[0])
    vn.running_mean_sq[:] = float(stats[1])
    vn.debiasing_term.fill_(float(stats[2]))
    return vn


def test_gae_matches_reference():
    g = golden("gae")
    for case in ("vn", "raw"):
        vn = _vn_from(g["vn_stats"]) if case == "vn" else None
        adv, ret, _ = po.compute_returns(g[case + "_rewards"], g[case + "_values_in"], g[case + "_masks"],
                                         g[case + "_next"], vn)
        np.testing.assert_array_equal(adv, g[case + "_adv"])
        np.testing.assert_array_equal(ret[:-1], g[case + "_ret"][:-1])


def test_ppo_loss_and_gradients_match_reference():
    g = golden("ppo_L11")
    L = int(g["L"])
    sd = mo.make_weights(int(g["wseed"]), L, profile="trained")
    for p in sd.values():
        p.requires_grad_(True)
    vn = _vn_from(g["vn_stats_before"])
    loss, vl, pl, ent, ratio = po.ppo_loss(sd, L, g["obs"], g["actions"], g["ava"], g["old_logp"], g["adv"],
                                           g["value_preds"], g["returns"], g["active"], vn)
    loss.backward()
    assert abs(vl.item() - float(g["value_loss"])) < 1e-5 * max(1, abs(float(g["value_loss"])))
    assert abs(pl.item() - float(g["policy_loss"])) < 1e-5
    assert abs(ent.item() - float(g["entropy"])) < 1e-5
    assert abs(ratio.mean().item() - float(g["ratio_mean"])) < 1e-5
    proj = np.random.RandomState(99)
    sq = 0.0
    for name, st in zip(g["grad_names"], g["grad_stats"]):
        p = sd[str(name)]
        gr = np.zeros(p.shape, np.float32) if p.grad is None else p.grad.numpy()
        r = proj.standard_normal(gr.shape)
        assert (p.grad is not None) == bool(st[3]), name
        scale = max(st[1], 1e-5 * float(g["grad_norm"]))   # key biases have zero gradient (softmax shift invariance)
        assert abs(gr.sum(dtype=np.float64) - st[0]) < 2e-4 * scale * np.sqrt(gr.size), name
        assert abs(np.sqrt((gr.astype(np.float64) ** 2).sum()) - st[1]) < 1e-4 * scale, name
        assert abs((gr * r).sum(dtype=np.float64) - st[2]) < 2e-4 * scale * np.sqrt(gr.size), name
        if "grad/" + str(name) in g.files:
            np.testing.assert_allclose(gr, g["grad/" + str(name)], atol=1e-4 * scale, rtol=1e-4)
        sq += float((gr.astype(np.float64) ** 2).sum())
    assert abs(np.sqrt(sq) - float(g["grad_norm"])) < 1e-4 * float(g["grad_norm"])


@pytest.mark.parametrize("tag,kw", [("fx", dict(fixed=True)), ("sa", dict(standalone=True))])
def test_env_variants_match_reference(tag, kw):
    """Env(fixed=True) (ENV:58-62) and step(standalone=True) (ENV:81-92), each env of the fixture stepped directly."""
    g = golden("env_W16_variants")
    W, B, steps = int(g["W"]), int(g["B"]), int(g["steps"])
    env = eo.DcmlEnvOracle(B, W, g["base"], seed=int(g["seed"]), double_reset=False, **kw)
    obs, sob, ava = env.reset()
    np.testing.assert_allclose(obs, g[tag + "_obs"][:, 0], rtol=1e-12, atol=1e-14)
    np.testing.assert_array_equal(ava, g[tag + "_ava"][:, 0])
    for t in range(1, steps + 1):
        obs, sob, rew, dones, infos, ava = env.step(g[tag + "_act"][:, t - 1])
        np.testing.assert_allclose([i[0]["delay"] for i in infos], g[tag + "_delay"][:, t - 1], rtol=1e-12)
        np.testing.assert_allclose([i[0]["payment"] for i in infos], g[tag + "_pay"][:, t - 1], rtol=1e-12)
        np.testing.assert_allclose(rew[:, 0, 0], g[tag + "_rew"][:, t - 1], rtol=1e-12)
        np.testing.assert_array_equal(dones[:, 0], g[tag + "_done"][:, t - 1])
        np.testing.assert_allclose(obs, g[tag + "_obs"][:, t], rtol=1e-12, atol=1e-14)
        np.testing.assert_allclose(sob[:, 0], g[tag + "_sob"][:, t], rtol=1e-12, atol=1e-14)
        np.testing.assert_array_equal(ava, g[tag + "_ava"][:, t])

# Random123 (Salmon et al., SC'11) known-answer vectors of Philox4x32-10: (counter[4], key[2]) -> output[4]
PHILOX_KAT = (
    ((0x00000000, 0x00000000, 0x00000000, 0x00000000), (0x00000000, 0x00000000),
     (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    ((0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff), (0xffffffff, 0xffffffff),
     (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
     (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
)


def test_philox_known_answers():
    """oracle/philox.py is the Philox4x32-10 of Random123 (the named draws of DESIGN.md section 4 rest on it)."""
    for ctr, key, want in PHILOX_KAT:
        got = px.philox4x32(*ctr, seed=key[0] | (key[1] << 32))
        assert tuple(int(x) for x in got) == want
    # vectorised over counters: same words as the scalar calls
    c0 = np.arange(5, dtype=np.uint32)
    got = px.philox4x32(c0, 7, 9, 11, seed=0x0123456789ABCDEF)
    for i in range(5):
        one = px.philox4x32(int(c0[i]), 7, 9, 11, seed=0x0123456789ABCDEF)
        assert [int(g[i]) for g in got] == [int(x) for x in one]
