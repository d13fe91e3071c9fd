#!/usr/bin/env python
"""Generate the golden fixtures in this directory by running the UNMODIFIED reference
(/root/reference, brithbutter/mat-dcml) in the build container.

    python tests/golden/make_golden.py            # regenerates every *.npz here

The reference ships no tests or golden vectors, so these fixtures are
what pins the oracle: ``tests/test_oracle_golden.py`` checks ``oracle/`` against them on CPU,
``tests/test_*_gpu.py`` check the CUDA path against them on the GPU box (where
``/root/reference`` does not exist).  Nothing is copied from the reference: it is imported
from where it lies, its RNG entry points (``random`` / ``np.random`` module globals of the
four simulator modules) are replaced by shims that return the named Philox draws of
``oracle/philox.py``, and only inputs and outputs are stored.

Each family runs in a subprocess because the simulator star-imports its constants
(``DCML_BID_FIRST_MA_ENV_SingleProcess.py:4``) and W != 100 needs them patched before import.
"""
import argparse
import math
import os
import subprocess
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.dont_write_bytecode = True


def _ref_paths():
    sys.path[:0] = [REF, os.path.join(REF, "mat_src")]
    stub = types.ModuleType("tensorboardX")
    stub.SummaryWriter = type("SummaryWriter", (), {"__init__": lambda self, *a, **k: None})
    sys.modules["tensorboardX"] = stub


def _args(**over):
    from mat.config import get_config
    a = get_config().parse_known_args([])[0]
    a.algorithm_name, a.env_name = "mat", "DCML"
    for k, v in over.items():
        setattr(a, k, v)
    return a


# ------------------------------------------------------------------------------------------
# policy
# ------------------------------------------------------------------------------------------
def synth_policy_inputs(B, L, seed):
    """Env-like observations: features in [0,1.1], ~40% workers disabled (ava [1,0])."""
    rng = np.random.RandomState(seed)
    obs = rng.uniform(0.0, 1.1, size=(B, L, 7))
    dis = rng.uniform(size=(B, L)) < 0.4
    dis[:, -1] = False
    obs[dis, 2:6] = 1.0
    ava = np.ones((B, L, 2), np.int64)
    ava[dis, 1] = 0
    return obs.reshape(B * L, 7), ava.reshape(B * L, 2)


def gen_policy(out, L, B, profile, wseed, semi=True):
    _ref_paths()
    import torch
    from mat.algorithms.mat.algorithm.transformer_policy import TransformerPolicy
    from DCML_ENVs.DCML_utils.DCML_ActionSpace import Action_Space
    from oracle import mat_oracle as mo

    W = L - 1
    space = Action_Space(2, semi_index=-1 if semi else 0, extra=True)
    pol = TransformerPolicy(_args(), [7], [W + 2], space, L)
    assert pol.action_type == ("Semi_Discrete" if semi else "Discrete")
    sd = mo.make_weights(wseed, L, profile=profile, semi=semi)
    pol.transformer.load_state_dict(mo.full_state_dict(sd, L), strict=True)
    pol.eval()
    obs, ava = synth_policy_inputs(B, L, 100 + wseed)
    cent = np.zeros((B * L, W + 2))
    res = {"obs": obs, "ava": ava, "L": L, "B": B, "wseed": wseed, "profile": profile, "semi": int(semi)}
    with torch.no_grad():
        res["values"] = pol.get_values(cent, obs, None, None, ava).numpy()
        strides = [1, 2, 10, L] if semi else [4]
        for s in strides:
            v, a, lp, _, _ = pol.get_actions(cent, obs, None, None, None, ava.copy(), deterministic=True, stride=s)
            res["det%d_values" % s], res["det%d_actions" % s], res["det%d_logp" % s] = v.numpy(), a.numpy(), lp.numpy()
        # stochastic sample from the reference's own RNG, to be re-scored by evaluate_actions
        torch.manual_seed(7)
        v, a, lp, _, _ = pol.get_actions(cent, obs, None, None, None, ava.copy(), deterministic=False)
        res["smp_actions"], res["smp_logp"] = a.numpy(), lp.numpy()
    active = np.ones((B * L, 1), np.float32)
    active[::5] = 0.0
    for tagname, acts in (("det", res["det1_actions" if semi else "det4_actions"]), ("smp", res["smp_actions"])):
        acts = acts.astype(np.float32)
        lp_t, v_t, ent_t = pol.transformer(cent.reshape(B, L, -1), obs.reshape(B, L, 7), acts.reshape(B, L, 1),
                                           ava.reshape(B, L, 2).copy())
        v2, lp2, ent_s = pol.evaluate_actions(cent, obs, None, None, acts, None, ava.copy(), torch.from_numpy(active))
        res["ev_%s_logp" % tagname] = lp_t.detach().numpy().reshape(-1, 1)
        res["ev_%s_entropy" % tagname] = ent_t.detach().numpy().reshape(-1, 1)
        res["ev_%s_values" % tagname] = v2.detach().numpy()
        res["ev_%s_entropy_scalar" % tagname] = ent_s.detach().numpy()
    res["active"] = active
    np.savez_compressed(out, **res)
    print("wrote", out)


# ------------------------------------------------------------------------------------------
# env (reference Env + ShareDummyVecEnv with injected named draws)
# ------------------------------------------------------------------------------------------
class _Ctx:
    pass


def gen_env(out, W, B, steps, seed, real_workloads, shannon=False, variants=False, single=False):
    _ref_paths()
    import DCML_ENVs.DCML_utils.DCML_Config as cfg
    if W != 100 or shannon:                        # constants are star-imported
        cfg.WORKER_NUMBER_MAX = W
        cfg.SOB_DIM = 2 * W + 2 if shannon else W + 2
    from oracle import philox as px
    from oracle import env_oracle as eo

    if real_workloads:
        os.chdir(REF)
        base = []
        with open("data/workloads.txt", "rb") as f:
            for _ in range(W):
                base.append(np.load(f, allow_pickle=True))
        base = np.asarray(base, np.float64)
    else:
        base = eo.synthetic_workloads(W, seed=11)
        tmp = tempfile.mkdtemp()
        os.makedirs(os.path.join(tmp, "data"))
        with open(os.path.join(tmp, "data", "workloads.txt"), "wb") as f:
            for row in base:
                np.save(f, row)
        os.chdir(tmp)

    import DCML_Master
    import DCML_Worker_TIMESLOT_MultiProcess as WK
    import DCML_ENVs.DCML_Basic_Env as BE
    import DCML_BID_FIRST_MA_ENV_SingleProcess as ENV
    from mat.envs.env_wrappers import ShareDummyVecEnv

    ctx = _Ctx()
    ctx.t, ctx.e, ctx.w, ctx.k, ctx.in_reset = 0, 0, 0, 0, False
    ctx.r = {}
    ctx.draws = {}
    kmax = int(math.floor(0.8 * W))

    def draw():
        key = (ctx.e, ctx.t)
        if key not in ctx.draws:
            ctx.draws[key] = eo._Draws(seed, ctx.e, ctx.t, W)
            ctx.draws[key]._jit = ctx.draws[key].jitter()
        return ctx.draws[key]

    class PyRandom:
        @staticmethod
        def uniform(a, b):
            if (a, b) == (50, 60):                  # master transmit power, DCML_Master.py:42
                pm, _, _ = draw().shannon_block(ctx.r.get(ctx.e, 0))
                return a + (b - a) * pm
            if b == 1:                              # geometric retry draw, Worker.process
                v = draw().retry(ctx.w, ctx.k)
                ctx.k += 1
                return float(v)
            return 0.5 * (a + b)                    # master Pr: never observed (HETEROGENEOUS)

        @staticmethod
        def random():
            if ctx.in_reset:
                return 0.5                          # update_local_workload_delay: value unused
            return draw().done_coin()

        @staticmethod
        def randint(a, b):
            x = draw().reset_scalars(ctx.r.get(ctx.e, 0))
            if (a, b) == (1, 80):
                return 1 + int(px.bounded(x[0], kmax))
            if (a, b) == (0, 19):
                return int(px.bounded(x[1], 20))
            if (a, b) == (eo.R_MIN, eo.R_HI):
                return eo.R_MIN + int(px.bounded(x[2], eo.R_HI - eo.R_MIN + 1))
            if (a, b) == (eo.C_MIN, eo.C_HI):
                return eo.C_MIN + int(px.bounded(x[3], eo.C_HI - eo.C_MIN + 1))
            return a                                # Master.__init__: overwritten by the first reset

        @staticmethod
        def seed(*a):
            pass

    class NpRandom:
        @staticmethod
        def uniform(lo, hi, size=None):
            if (lo, hi) == (0.8, 1.2):
                return lo + (hi - lo) * draw()._jit[ctx.w]
            if (lo, hi) == (0, 0.95):
                u, _ = draw().worker_block(ctx.r.get(ctx.e, 0))
                return lo + (hi - lo) * u
            if (lo, hi) == (10, 20):                # worker transmit power, DCML_Basic_Env.py:20
                return lo + (hi - lo) * draw().shannon_block(ctx.r.get(ctx.e, 0))[1]
            if (lo, hi) == (10, 100):               # distances, DCML_Master.py:43
                return lo + (hi - lo) * draw().shannon_block(ctx.r.get(ctx.e, 0))[2]
            raise AssertionError(("unexpected np.random.uniform", lo, hi, size))

        @staticmethod
        def poisson(lam=1.0, size=None):
            return np.full(size, 3)                 # feeds Worker.price only (DYNAMIC_PRICE False)

        @staticmethod
        def choice(n, k, replace=True):
            assert n == W and not replace
            _, key = draw().worker_block(ctx.r.get(ctx.e, 0))
            return np.lexsort((np.arange(W), key))[:k]

    class NpProxy:
        random = NpRandom

        def __getattr__(self, name):
            return getattr(np, name)

    for mod in (DCML_Master, WK, BE, ENV):
        if hasattr(mod, "random"):
            mod.random = PyRandom
        if hasattr(mod, "np"):
            mod.np = NpProxy()

    w_init, w_process, e_reset, e_step = WK.Worker.__init__, WK.Worker.process, ENV.Env.reset, ENV.Env.step

    def worker_init(self, *a, **k):
        self._idx = ctx.w
        w_init(self, *a, **k)
        ctx.w += 1

    def worker_process(self, *a, **k):
        ctx.w, ctx.k = self._idx, 0
        return w_process(self, *a, **k)

    def env_reset(self, *a, **k):
        ctx.e, ctx.in_reset = self._e, True
        try:
            return e_reset(self, *a, **k)
        finally:
            ctx.in_reset = False
            ctx.r[self._e] = ctx.r.get(self._e, 0) + 1

    def env_step(self, *a, **k):
        ctx.e = self._e
        ctx.r[self._e] = 0
        return e_step(self, *a, **k)

    WK.Worker.__init__, WK.Worker.process = worker_init, worker_process
    ENV.Env.reset, ENV.Env.step = env_reset, env_step

    def make(e, **kw):
        def fn():
            ctx.e, ctx.w = e, 0
            env = ENV.Env(**kw)
            env._e = e
            return env
        return fn

    if variants:
        # Env(fixed=True) (ENV:58-62: actions ignored, all available workers, K = floor(0.7 N)) and step(standalone=True)
        # (ENV:81-92: uncoded on worker 0, reward x 1.5), each env stepped directly (no vec wrapper, no second reset)
        L = W + 1
        res = {"W": W, "B": B, "steps": steps, "seed": seed, "base": base}
        rng = np.random.RandomState(seed + 5)
        for tag, kw, skw in (("fx", dict(fixed=True), {}), ("sa", {}, dict(standalone=True))):
            ctx.draws.clear(); ctx.r.clear()
            acc = {}
            for e in range(B):
                ctx.t = 0
                env = make(e, **kw)()
                ctx.r[e] = 0
                obs, sob, ava = env.reset()
                rows = {"obs": [obs], "sob": [sob[0]], "ava": [ava], "act": [], "rew": [], "delay": [], "pay": [], "done": []}
                for t in range(1, steps + 1):
                    ctx.t = t
                    sel = (rng.uniform(size=W) < 0.6).astype(np.float32) * ava[:W, 1]
                    act = np.concatenate([sel, [np.float32(rng.uniform(0.2, 1.0))]]).astype(np.float32)
                    obs, sob, rew, dones, info, ava = env.step(act.reshape(L, 1), **skw)
                    for k, v in zip(("obs", "sob", "ava", "act", "rew", "delay", "pay", "done"),
                                    (obs, sob[0], ava, act, rew[0, 0], info[0]["delay"], info[0]["payment"], dones[0])):
                        rows[k].append(v)
                for k, v in rows.items():
                    acc.setdefault(k, []).append(np.asarray(v))
            res.update({tag + "_" + k: np.asarray(v) for k, v in acc.items()})      # [env][step ...]
        np.savez_compressed(out, **res)
        print("wrote", out)
        return

    if single:
        # Env(multi_agent=False) (ENV:42-47,272-274; BASE:14-17) with a first reset(binary=True) (ENV:164-172): flat observation,
        # ONE share_obs row (64 + W wide after the binary reset, W + 2 after a step), one reward / done per env
        L = W + 1
        res = {"W": W, "B": B, "steps": steps, "seed": seed, "base": base}
        rng = np.random.RandomState(seed + 5)
        acc = {}
        for e in range(B):
            ctx.t = 0
            env = make(e, multi_agent=False)()
            ctx.r[e] = 0
            assert env.observation_space == [[7 * L]] and env.action_space[0].mixed and env.action_space[0].high == W
            obs, sob, ava = env.reset(binary=True)
            assert obs.shape == (7 * L,) and sob.shape == (64 + W,) and ava.shape == (L, 2)
            rows = {"obs": [], "sob": [], "ava": [ava], "act": [], "rew": [], "delay": [], "pay": [], "done": []}
            acc.setdefault("obs0", []).append(obs); acc.setdefault("sob0", []).append(sob)
            for t in range(1, steps + 1):
                ctx.t = t
                sel = (rng.uniform(size=W) < 0.6).astype(np.float32) * ava[:W, 1]
                act = np.concatenate([sel, [np.float32(rng.uniform(0.2, 1.0))]]).astype(np.float32)
                obs, sob, rew, dones, info, ava = env.step(act.reshape(L, 1))
                assert obs.shape == (7 * L,) and sob.shape == (W + 2,) and rew.shape == (1,) and dones.shape == (1,)
                for k, v in zip(("obs", "sob", "ava", "act", "rew", "delay", "pay", "done"),
                                (obs, sob, ava, act, rew[0], info[0]["delay"], info[0]["payment"], dones[0])):
                    rows[k].append(v)
            for k, v in rows.items():
                acc.setdefault(k, []).append(np.asarray(v))
        res.update({k: np.asarray(v) for k, v in acc.items()})          # [env][step ...]
        np.savez_compressed(out, **res)
        print("wrote", out)
        return

    if shannon:
        # Shannon links (dormant in the shipped config): only an explicit reset(shannon_enable=True) and the
        # step that follows it see Shannon rates (Env.step re-resets without the flag, ENV:139).
        L = W + 1
        res = {"W": W, "B": B, "seed": seed, "base": base}
        rng = np.random.RandomState(seed + 5)
        keys = ("obs0", "sob0", "ava0", "act1", "rew1", "delay1", "pay1", "done1")
        acc = {k: [] for k in keys}
        for e in range(B):
            ctx.t = 0                               # construction jitter and the explicit reset are step-0 draws
            env = make(e)()
            ctx.r[e] = 0
            obs, sob, ava = env.reset(shannon_enable=True)
            assert sob.shape == (L, 2 * W + 2)
            sel = (rng.uniform(size=W) < 0.6).astype(np.float32) * ava[:W, 1]
            act = np.concatenate([sel, [np.float32(rng.uniform(0.2, 1.0))]]).astype(np.float32)
            ctx.t = 1
            o2, s2, rew, dones, info, a2 = env.step(act.reshape(L, 1))
            for k, v in zip(keys, (obs, sob[0], ava, act, rew[0, 0], info[0]["delay"], info[0]["payment"], dones[0])):
                acc[k].append(v)
        res.update({k: np.asarray(v) for k, v in acc.items()})
        np.savez_compressed(out, **res)
        print("wrote", out)
        return
    vec = ShareDummyVecEnv([make(e) for e in range(B)])
    L = W + 1
    res = {"W": W, "B": B, "steps": steps, "seed": seed, "base": base}
    obs, sob, ava = vec.reset()
    assert obs.shape == (B, L, 7) and sob.shape == (B, L, W + 2) and ava.shape == (B, L, 2)
    res["obs0"], res["sob0"], res["ava0"] = obs, sob[:, 0, :], ava
    assert np.all(sob == sob[:, :1, :])
    rng = np.random.RandomState(seed + 5)
    for t in range(1, steps + 1):
        ctx.t = t
        sel = (rng.uniform(size=(B, W)) < rng.uniform(0.2, 0.9, size=(B, 1))).astype(np.float32)
        sel = sel * ava[:, :W, 1]                   # the policy never selects a disabled worker
        ratio = rng.uniform(0.05, 1.2, size=(B, 1)).astype(np.float32)
        if t == 2:
            sel[0] = 0.0                            # the N == 0 branch (ENV:81-92)
        if t == 3:
            ratio[0] = -0.3                         # K clamped up to 1
        act = np.concatenate([sel, ratio], axis=1).astype(np.float32)[:, :, None]
        obs, sob, rew, dones, infos, ava = vec.step(act)
        assert np.all(sob == sob[:, :1, :]) and np.all(rew == rew[:, :1, :]) and np.all(dones == dones[:, :1])
        res["act%d" % t] = act
        res["obs%d" % t], res["sob%d" % t], res["ava%d" % t] = obs, sob[:, 0, :], ava
        res["rew%d" % t], res["done%d" % t] = rew[:, 0, 0], dones[:, 0]
        res["delay%d" % t] = np.array([i[0]["delay"] for i in infos], np.float64)
        res["pay%d" % t] = np.array([i[0]["payment"] for i in infos], np.float64)
    np.savez_compressed(out, **res)
    print("wrote", out)


# ------------------------------------------------------------------------------------------
# GAE + PPO update
# ------------------------------------------------------------------------------------------
def gen_gae(out):
    _ref_paths()
    import torch
    from mat.utils.shared_buffer import SharedReplayBuffer
    from mat.utils.valuenorm import ValueNorm
    from DCML_ENVs.DCML_utils.DCML_ActionSpace import Action_Space

    T, N, L = 12, 5, 7
    rng = np.random.RandomState(3)
    res = {}
    for case, use_vn in (("vn", True), ("raw", False)):
        args = _args(episode_length=T, n_rollout_threads=N, use_valuenorm=use_vn, use_popart=False)
        buf = SharedReplayBuffer(args, L, [7], [L + 1], Action_Space(2, semi_index=-1, extra=True), "DCML")
        buf.rewards[:] = rng.standard_normal(buf.rewards.shape) * 50 - 200
        buf.value_preds[:] = rng.standard_normal(buf.value_preds.shape)
        buf.masks[:] = (rng.uniform(size=(T + 1, N, 1, 1)) > 0.8)
        nxt = rng.standard_normal((N, L, 1)).astype(np.float32)
        vn = None
        if use_vn:
            vn = ValueNorm(1)
            for _ in range(3):
                vn.update(torch.from_numpy((rng.standard_normal((64, 1)) * 40 - 150).astype(np.float32)))
            res["vn_stats"] = np.array([vn.running_mean.item(), vn.running_mean_sq.item(), vn.debiasing_term.item()],
                                       np.float32)
        res[case + "_rewards"], res[case + "_values_in"] = buf.rewards.copy(), buf.value_preds.copy()
        res[case + "_masks"], res[case + "_next"] = buf.masks.copy(), nxt
        buf.compute_returns(nxt, vn)
        res[case + "_adv"], res[case + "_ret"] = buf.advantages.copy(), buf.returns.copy()
    np.savez_compressed(out, **res)
    print("wrote", out)


def gen_ppo(out, L, M, wseed):
    """One MATTrainer.ppo_update on the reference with lr=0 and no effective grad clipping:
    stores the losses and per-tensor gradient statistics (full gradients of tensors <= 128 el.)."""
    _ref_paths()
    import torch
    from mat.algorithms.mat.algorithm.transformer_policy import TransformerPolicy
    from mat.algorithms.mat.mat_trainer import MATTrainer
    from DCML_ENVs.DCML_utils.DCML_ActionSpace import Action_Space
    from oracle import mat_oracle as mo

    W = L - 1
    args = _args(lr=0.0, use_valuenorm=True, max_grad_norm=1e9, n_rollout_threads=1)
    pol = TransformerPolicy(args, [7], [W + 2], Action_Space(2, semi_index=-1, extra=True), L)
    sd = mo.make_weights(wseed, L, profile="trained")
    pol.transformer.load_state_dict(mo.full_state_dict(sd, L), strict=True)
    tr = MATTrainer(args, pol, L)
    obs, ava = synth_policy_inputs(M, L, 200 + wseed)
    rng = np.random.RandomState(17)
    cent = np.zeros((M * L, W + 2), np.float32)
    with torch.no_grad():
        torch.manual_seed(3)
        v, a, lp, _, _ = pol.get_actions(cent, obs, None, None, None, ava.copy(), deterministic=False)
    actions = a.numpy().astype(np.float32)
    old_lp = (lp.numpy() + 0.3 * rng.standard_normal(lp.shape)).astype(np.float32)
    adv = rng.standard_normal((M * L, 1)).astype(np.float32)
    vpred = (v.numpy() + 0.3 * rng.standard_normal(v.shape)).astype(np.float32)
    ret = (rng.standard_normal((M * L, 1)) * 30 - 100).astype(np.float32)
    active = np.ones((M * L, 1), np.float32)
    active[::7] = 0.0
    for _ in range(2):
        tr.value_normalizer.update(torch.from_numpy((rng.standard_normal((64, 1)) * 30 - 100).astype(np.float32)))
    vn0 = np.array([tr.value_normalizer.running_mean.item(), tr.value_normalizer.running_mean_sq.item(),
                    tr.value_normalizer.debiasing_term.item()], np.float32)
    sample = (cent, obs.astype(np.float32), None, None, actions, vpred, ret, None, active, old_lp, adv,
              ava.astype(np.float32))
    pol.train()
    vl, gn, pl, ent, _, ratio = tr.ppo_update(sample)
    res = {"L": L, "M": M, "wseed": wseed, "obs": obs, "ava": ava, "actions": actions, "old_logp": old_lp,
           "adv": adv, "value_preds": vpred, "returns": ret, "active": active, "vn_stats_before": vn0,
           "value_loss": vl.item(), "policy_loss": pl.item(), "entropy": ent.item(), "grad_norm": float(gn),
           "ratio_mean": ratio.mean().item()}
    proj = np.random.RandomState(99)
    names, stats = [], []
    for name, p in pol.transformer.named_parameters():
        g = np.zeros(p.shape, np.float32) if p.grad is None else p.grad.numpy()
        r = proj.standard_normal(g.shape)
        names.append(name)
        stats.append([g.sum(dtype=np.float64), np.sqrt((g.astype(np.float64) ** 2).sum()), (g * r).sum(dtype=np.float64),
                      0.0 if p.grad is None else 1.0])
        if g.size <= 128:
            res["grad/" + name] = g
    res["grad_names"], res["grad_stats"] = np.array(names), np.array(stats)
    np.savez_compressed(out, **res)
    print("wrote", out)


def gen_buffer(out):
    """SharedReplayBuffer.insert / after_update / feed_forward_generator_transformer of the reference on random content
    with an injected permutation (``torch.randperm`` replaced for the call): the 12-tuples of three minibatches."""
    _ref_paths()
    import torch
    from mat.utils.shared_buffer import SharedReplayBuffer
    from DCML_ENVs.DCML_utils.DCML_ActionSpace import Action_Space
    import mat.utils.shared_buffer as SB

    T, N, L, W = 6, 4, 5, 4
    rng = np.random.RandomState(23)
    args = _args(episode_length=T, n_rollout_threads=N, use_valuenorm=False, use_popart=False)
    buf = SharedReplayBuffer(args, L, [7], [W + 2], Action_Space(2, semi_index=-1, extra=True), "DCML")
    f32 = lambda *sh: rng.standard_normal(sh).astype(np.float32)
    res = {"T": T, "N": N, "L": L, "W": W}
    first = {"share_obs": f32(N, L, W + 2), "obs": f32(N, L, 7), "ava": (rng.uniform(size=(N, L, 2)) > 0.3).astype(np.float32)}
    first["share_obs"][:] = first["share_obs"][:, :1]                 # the env replicates one row over the agents
    buf.share_obs[0], buf.obs[0], buf.available_actions[0] = first["share_obs"], first["obs"], first["ava"]
    res.update({"first_" + k: v for k, v in first.items()})
    for t in range(T):
        so = f32(N, 1, W + 2).repeat(L, axis=1)
        step = dict(share_obs=so, obs=f32(N, L, 7), actions=f32(N, L, 1), logp=f32(N, L, 1), values=f32(N, L, 1),
                    rewards=f32(N, 1, 1).repeat(L, axis=1), masks=(rng.uniform(size=(N, 1, 1)) > 0.3).astype(np.float32).repeat(L, axis=1),
                    ava=(rng.uniform(size=(N, L, 2)) > 0.3).astype(np.float32))
        buf.insert(step["share_obs"], step["obs"], np.zeros((N, L, 1, 64), np.float32), np.zeros((N, L, 1, 64), np.float32),
                   step["actions"], step["logp"], step["values"], step["rewards"], step["masks"], None, None, step["ava"])
        res.update({"s%d_%s" % (t, k): v for k, v in step.items()})
    buf.compute_returns(f32(N, L, 1), None)
    res["next_value"] = buf.value_preds[-1].copy()
    adv = f32(T, N, L, 1)
    perm = rng.permutation(T * N)
    res["adv"], res["perm"] = adv, perm.astype(np.int64)
    real = torch.randperm
    torch.randperm = lambda n, **k: torch.from_numpy(perm.copy())
    try:
        names = ("share_obs", "obs", "rnn", "rnn_c", "actions", "value_preds", "returns", "masks", "active", "old_logp", "adv", "ava")
        for i, sample in enumerate(buf.feed_forward_generator_transformer(adv, num_mini_batch=3)):
            for k, v in zip(names, sample):
                if not k.startswith("rnn"):
                    res["mb%d_%s" % (i, k)] = np.asarray(v)
    finally:
        torch.randperm = real
    buf.after_update()
    res["after_obs0"], res["after_share0"], res["after_masks0"], res["after_ava0"] = (
        buf.obs[0].copy(), buf.share_obs[0, :, 0].copy(), buf.masks[0].copy(), buf.available_actions[0].copy())
    np.savez_compressed(out, **res)
    print("wrote", out)


# ------------------------------------------------------------------------------------------
# preset benchmark samples (Env(preset=True) reader + modify_preset)
# ------------------------------------------------------------------------------------------
def gen_preset(out, rows=24):
    """What the reference's ``Env(preset=True)`` holds after reading ``data/dcml_benchmark/Sample_1*`` (the first
    ``rows`` episodes) and after the benchmark script's ``modify_preset`` calls (DCML_MAT_ALT_Benchmark.py:39,122)."""
    _ref_paths()
    os.chdir(REF)                                     # the env opens data/... relative to the cwd
    from DCML_BID_FIRST_MA_ENV_SingleProcess import Env
    env = Env(preset=True)
    res = {"n_rows": np.int64(len(env.master_status)), "master": env.master_status[:rows].astype(np.float64),
           "worker_prs": env.worker_Prs[:rows].copy(), "disable_rates": env.disable_rates[:rows].copy()}
    env.modify_preset(R=(2 ** 19), C=(9 * (2 ** 10) / 10))
    res["master_RC"] = env.master_status[:rows].astype(np.float64)
    env.modify_preset(disable_rate=3 * 8)
    res["disable_rates_24"] = env.disable_rates[:rows].copy()
    env.modify_preset(Pr=0.3)
    res["worker_prs_03"] = env.worker_Prs[:rows].copy()
    np.savez_compressed(out, **res)
    print("wrote", out)


FAMILIES = {
    "policy_L21_trained": lambda o: gen_policy(o, 21, 5, "trained", 3),
    "policy_L21_init": lambda o: gen_policy(o, 21, 5, "init", 4),
    "policy_L101_trained": lambda o: gen_policy(o, 101, 2, "trained", 5),
    "policy_L13_discrete": lambda o: gen_policy(o, 13, 4, "trained", 6, semi=False),
    "env_W100": lambda o: gen_env(o, 100, 3, 6, 1, True),
    "env_W20": lambda o: gen_env(o, 20, 6, 8, 2, False),
    # wider families (32 envs x 16 steps): several hundred reference env-steps through every floor / ceil / % branch of
    # Worker.process at the default worker count (real workloads.txt) and at the C3 bench shape (50 workers)
    "env_W100_wide": lambda o: gen_env(o, 100, 32, 16, 11, True),
    "env_W50_wide": lambda o: gen_env(o, 50, 32, 16, 12, False),
    "env_W12_shannon": lambda o: gen_env(o, 12, 8, 1, 4, False, shannon=True),
    "env_W16_variants": lambda o: gen_env(o, 16, 6, 5, 7, False, variants=True),
    "env_W16_single_binary": lambda o: gen_env(o, 16, 6, 4, 9, False, single=True),
    "gae": gen_gae,
    "buffer": gen_buffer,
    "ppo_L11": lambda o: gen_ppo(o, 11, 6, 8),
    "preset_sample1": gen_preset,
}

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("family", nargs="?")
    ns = ap.parse_args()
    if ns.family:
        FAMILIES[ns.family](os.path.join(HERE, ns.family + ".npz"))
    else:
        for fam in FAMILIES:
            subprocess.check_call([sys.executable, os.path.abspath(__file__), fam],
                                  env=dict(os.environ, PYTHONDONTWRITEBYTECODE="1"))
