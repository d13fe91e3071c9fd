"""The reference's deterministic benchmark mode on the device-resident path.

``Env(preset=True)`` reads one pair of sample files (``data/dcml_benchmark/Sample_{k}master_states.npy`` and
``Sample_{k}worker_states.npy``, written by ``Env.generate_preset_data``,
DCML_BID_FIRST_MA_ENV_SingleProcess.py:25-32, 309-342) and serves row ``i`` of them to the ``i``-th reset
(:174-178, 191-194); ``modify_preset`` overwrites whole columns (:344-353); ``DCML_MAT_ALT_Benchmark.py:103-140``
then runs 1000 deterministic stride-10 decisions per sweep setting and averages reward, completion time and
payment.  ``PresetSamples`` is that reader; ``run_benchmark`` is that loop with the preset rows spread over the
envs of a ``DcmlVecEnv`` (row ``r*N + b`` goes to env ``b`` in round ``r``), every decision batched on the GPU.
"""
import numpy as np
import torch


class PresetSamples:
    """Rows of (R, C, master Pr), per-worker transmission-failure probabilities and disable counts."""

    def __init__(self, master_states, worker_prs, disable_rates):
        self.master_status = np.array(master_states, dtype=np.float64)
        self.worker_Prs = np.array(worker_prs, dtype=np.float64)
        self.disable_rates = np.array(disable_rates, dtype=np.int64)
        n = len(self.master_status)
        if self.master_status.ndim != 2 or self.master_status.shape[1] != 3:
            raise ValueError("master states must be (episodes, 3) rows of [R, C, Pr]")
        if self.worker_Prs.ndim != 2 or len(self.worker_Prs) != n or self.disable_rates.shape != (n,):
            raise ValueError("worker states must hold (episodes, W) probabilities and (episodes,) disable counts")

    @classmethod
    def load(cls, master_path, worker_path):
        """…SingleProcess.py:25-30: the master file is one ``np.save`` record, the worker file two consecutive ones."""
        with open(master_path, "rb") as f:
            master = np.load(f, allow_pickle=True)
        with open(worker_path, "rb") as f:
            prs = np.load(f, allow_pickle=False)
            dis = np.load(f, allow_pickle=False)
        return cls(master, prs, dis)

    def save(self, master_path, worker_path):
        """The same two files (…SingleProcess.py:330-342)."""
        with open(master_path, "wb") as f:
            np.save(f, self.master_status)
        with open(worker_path, "wb") as f:
            np.save(f, self.worker_Prs)
            np.save(f, self.disable_rates)

    def __len__(self):
        return len(self.master_status)

    @property
    def n_workers(self):
        return self.worker_Prs.shape[1]

    def modify_preset(self, R=None, C=None, Pr=None, disable_rate=None):
        """…SingleProcess.py:344-353 (``Pr`` overwrites every worker's probability, not the master's)."""
        if R is not None:
            self.master_status[:, 0] = R
        if C is not None:
            self.master_status[:, 1] = C
        if disable_rate is not None:
            self.disable_rates[:] = disable_rate
        if Pr is not None:
            self.worker_Prs[:] = Pr
        return self

    def apply(self, env, start=0):
        """Give env ``b`` the context of row ``start + b`` (wrapping around) and return the new observation."""
        if env.W != self.n_workers:
            raise ValueError("samples hold %d workers, the env has %d" % (self.n_workers, env.W))
        idx = (start + np.arange(env.N)) % len(self)
        return env.set_context(R=self.master_status[idx, 0], C=self.master_status[idx, 1], worker_pr=self.worker_Prs[idx],
                               n_disabled=self.disable_rates[idx].astype(np.int32))


@torch.no_grad()
def run_benchmark(policy, env, samples, n_episodes=1000, stride=10, deterministic=True):
    """DCML_MAT_ALT_Benchmark.py:115-140 for ONE sweep setting: ``n_episodes`` preset contexts, one deterministic
    decision and one env step each.  Returns ``{"reward", "ct", "payment"}`` means plus the per-episode arrays
    (device tensors) and the mean device time of one batched decision in seconds."""
    N, L = env.N, env.L
    model = policy.transformer
    keep, model.output_device = model.output_device, env.device         # actions stay on the GPU for the env step
    try:
        return _run(policy, env, samples, n_episodes, stride, deterministic, N, L)
    finally:
        model.output_device = keep


def _run(policy, env, samples, n_episodes, stride, deterministic, N, L):
    env.reset_device()
    rew, ct, pay = [], [], []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_ms, rounds, done = 0.0, 0, 0
    while done < n_episodes:
        obs, _, ava = samples.apply(env, start=done)
        e0.record()
        _, actions, _, _, _ = policy.get_actions(None, obs.reshape(N * L, -1), None, None, None, ava.reshape(N * L, -1),
                                                 deterministic=deterministic, stride=stride)
        e1.record()
        _, _, rewards, _, delay, payment, _ = env.step_device(actions.reshape(N, L, 1))
        k = min(N, n_episodes - done)
        rew.append(rewards[:k].double().clone()); ct.append(delay[:k].double().clone()); pay.append(payment[:k].double().clone())
        e1.synchronize()
        t_ms += e0.elapsed_time(e1)
        done += k
        rounds += 1
    rew, ct, pay = torch.cat(rew), torch.cat(ct), torch.cat(pay)
    return {"reward": float(rew.mean()), "ct": float(ct.mean()), "payment": float(pay.mean()), "rewards": rew, "cts": ct,
            "payments": pay, "decision_s": t_ms * 1e-3 / max(rounds, 1)}


def sweep(policy, env, samples, settings, n_episodes=1000, stride=10):
    """The script's outer loop (DCML_MAT_ALT_Benchmark.py:111-140): one ``modify_preset(**setting)`` per entry, e.g.
    ``[{"disable_rate": 8 * i} for i in range(11)]``; returns the ``w_cts`` / ``w_payments`` columns it saves."""
    w_cts, w_payments, rewards = [], [], []
    for s in settings:
        cur = PresetSamples(samples.master_status, samples.worker_Prs, samples.disable_rates).modify_preset(**s)
        r = run_benchmark(policy, env, cur, n_episodes=n_episodes, stride=stride)
        w_cts.append([r["ct"]]); w_payments.append([r["payment"]]); rewards.append(r["reward"])
    return np.array(w_cts), np.array(w_payments), np.array(rewards)
