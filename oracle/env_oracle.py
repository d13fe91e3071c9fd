"""CPU restatement of the DCML simulator step/reset (test infrastructure; see oracle/__init__.py).

Float64 / Python-int arithmetic in the reference's own evaluation order, one
env and one worker at a time like the reference's single-process loop, so it
also serves as the timed CPU baseline.  All randomness comes from the named
Philox draws of ``oracle/philox.py`` instead of the ``random`` / ``np.random``
global streams.  Citations: ``/root/reference`` root files
``DCML_BID_FIRST_MA_ENV_SingleProcess.py`` (ENV), ``DCML_Worker_TIMESLOT_MultiProcess.py``
(WORKER), ``DCML_Master.py`` (MASTER), ``DCML_ENVs/DCML_Basic_Env.py`` (BASE),
``DCML_ENVs/DCML_utils/DCML_Config.py`` (CFG), ``mat_src/mat/envs/env_wrappers.py`` (VEC).
"""
import math

import numpy as np

from . import philox as px

PERIOD = 20                       # WORKER:10, CFG:2
DATA_RATE = 150 * 2 ** 20         # CFG:5  NON_SHANNON_DATA_RATE
FREQUENCY = 2e9                   # WORKER:16
BIT_TO_BYTE = 4                   # WORKER:12
PR_MAX = 0.95                     # CFG:22
DONE_P = 0.8                      # CFG:25 CONTINUE_PROBABILITY (used as P(done), ENV:142)
R_MIN, R_MAX, C_MIN, C_MAX = 2 ** 10, 2 ** 20, 2 ** 5, 2 ** 10     # MASTER:6-9
R_HI = round(R_MAX * 1.1)         # MASTER:47
C_HI = round(C_MAX * 1.1)         # MASTER:48
ALPHA, BETA = 99, 1               # DCML_ENV_Functions.py:15-16
NOISE = 10 ** (-50 / 10)           # Shannon.py:9
B_TOTAL = 100 * 10 ** 9            # DCML_Master.py:29
RETRY_CAP = 4096                  # guard for the geometric loops (Pr <= 0.95 -> P(n>4096) ~ 1e-91)
SLOT_CAP = 4096                   # guard for the compute loop


def synthetic_workloads(W, seed=1):
    """``workloads.txt``-shaped base profile: W rows x 20 slots, multiples of 0.05 in [0,1],
    row means about 0.5."""
    rng = np.random.RandomState(seed)
    levels = np.arange(21) * 0.05
    centre = rng.uniform(0.45, 0.6, size=(W, 1))
    raw = np.clip(centre + 0.28 * rng.standard_normal((W, PERIOD)), 0.0, 1.0)
    idx = np.rint(raw / 0.05).astype(np.int64)
    idx[:, 0] = np.minimum(idx[:, 0], 16)          # never an all-ones row (App. D #20)
    return levels[idx]


class _Draws:
    """Named draws for one (env, step), generated in bulk per stream."""

    def __init__(self, seed, env, step, W):
        self.seed, self.env, self.step, self.W = seed, env, step, W
        self._retry = None
        self._nblk = 0

    def _grow_retry(self, nblk):
        w = np.arange(self.W, dtype=np.uint64)[:, None]
        b = np.arange(self._nblk, nblk, dtype=np.uint64)[None, :]
        x = px.philox4x32(self.env, self.step, (np.uint64(px.S_RETRY) << np.uint64(16)) | w, b, self.seed)
        new = np.stack([px.u01(v) for v in x], axis=-1).reshape(self.W, -1)
        self._retry = new if self._retry is None else np.concatenate([self._retry, new], axis=1)
        self._nblk = nblk

    def retry(self, w, n):
        if n >= 4 * self._nblk:
            self._grow_retry(max(16, 2 * self._nblk, n // 4 + 1))
        return self._retry[w, n]

    def jitter(self):
        """(W,20) uniforms in [0,1)."""
        w = np.arange(self.W, dtype=np.uint64)[:, None]
        b = np.arange(5, dtype=np.uint64)[None, :]
        x = px.philox4x32(self.env, self.step, (np.uint64(px.S_JITTER) << np.uint64(16)) | w, b, self.seed)
        return np.stack([px.u01(v) for v in x], axis=-1).reshape(self.W, PERIOD)

    def reset_scalars(self, r):
        return px.block(self.seed, self.env, self.step, px.S_RESET, 0, 16 * r)

    def done_coin(self):
        return float(px.u01(px.block(self.seed, self.env, self.step, px.S_RESET, 0, 1)[1]))

    def worker_block(self, r):
        w = np.arange(self.W, dtype=np.uint64)
        x = px.philox4x32(self.env, self.step, (np.uint64(px.S_WORKER) << np.uint64(16)) | w, r, self.seed)
        return px.u01(x[0]), x[1].astype(np.int64)

    def shannon_block(self, r):
        """-> (master power uniform, worker power uniforms (W,), distance uniforms (W,))"""
        w = np.arange(self.W, dtype=np.uint64)
        x = px.philox4x32(self.env, self.step, (np.uint64(px.S_WORKER) << np.uint64(16)) | w, r, self.seed)
        pm = float(px.u01(px.block(self.seed, self.env, self.step, px.S_RESET, 0, 16 * r + 2)[0]))
        return pm, px.u01(x[2]), px.u01(x[3])


def jitter_workload(base, u):
    """WORKER:39,111  clip(base * uniform(0.8, 1.2), 0, 1); numpy's uniform is low + (high-low)*u."""
    return np.clip(base * (0.8 + (1.2 - 0.8) * u), 0, 1)


def shannon_rates(n_workers, master_power, worker_power, distance):
    """Master.get_transmission_rate / Shannon.upload, .download (DCML_Master.py:41-45, Shannon.py:14-21)."""
    B = B_TOTAL / n_workers
    down = B * (np.log2(1 + (master_power * (distance ** -4)) / NOISE))
    up = B * (np.log2(1 + (worker_power * (distance ** -4)) / NOISE))
    return up, down


def process(lw, Pr, r, c, arrive, draw, w, DATA_RATE=DATA_RATE):
    """WORKER:46-112.  -> (n_retry, delay, prices list).  ``lw`` is the worker's 20-slot local
    workload (``local_workload`` is two copies of it, WORKER:131-133, indexed modulo 20)."""
    cost = math.ceil((9 * r - 3) * c) / FREQUENCY                                # WORKER:49-50
    n = 1
    k = 0
    while draw.retry(w, k) < Pr and n < RETRY_CAP:                               # WORKER:54-56
        n += 1
        k += 1
    k += 1
    t_down = ((math.ceil((r + 1) * c) * 1 * BIT_TO_BYTE) / DATA_RATE + 0.001) * n   # WORKER:60
    price = math.floor(t_down) * 0.1                                             # WORKER:65
    slot = int(math.floor(t_down + arrive))                                      # WORKER:66
    tp = slot % PERIOD                                                           # WORKER:67-69
    finish = slot
    avail = 0
    prices = []
    frac = t_down % 1
    if frac > lw[tp]:                                                            # WORKER:85-86
        cost += frac - lw[tp]
    t_up = 0.0
    while avail < cost and len(prices) < SLOT_CAP:                               # WORKER:87-106
        free = 1 - lw[tp]
        price += free * 1
        prices.append(price)
        avail += free
        tp += 1
        finish += 1
        if tp > PERIOD - 1:
            tp -= PERIOD
        while draw.retry(w, k) < Pr and n < RETRY_CAP:
            n += 1
            k += 1
        k += 1
        t_up = ((math.ceil(r) * 1 * BIT_TO_BYTE) / DATA_RATE + 0.001) * n + 0.02
    delay = finish - arrive - (avail - cost) + t_up                              # WORKER:108
    return n, delay, prices


class DcmlEnvOracle:
    """B independent DCML envs behind the reference's vec-env contract (VEC:713-744):
    ``reset() -> (obs, share_obs, ava)``, ``step(actions) -> (obs, share_obs, rews, dones, infos, ava)``.
    ``double_reset`` reproduces ShareDummyVecEnv's second reset on done (VEC:730-736)."""

    def __init__(self, n_envs, n_workers, base_workload=None, seed=1, double_reset=True, env0=0, shannon=False,
                 fixed=False, standalone=False):
        self.B, self.W, self.L = n_envs, n_workers, n_workers + 1
        self.shannon = shannon
        self.fixed, self.standalone = fixed, standalone       # Env(fixed=True) ENV:58-62; step(standalone=True) ENV:81
        self.rate_down = np.full((n_envs, n_workers), float(DATA_RATE))     # BASE:28-33 NON_SHANNON_DATA_RATE
        self.rate_up = np.full((n_envs, n_workers), float(DATA_RATE))
        self.seed, self.env0, self.double_reset = seed, env0, double_reset
        self.base = synthetic_workloads(n_workers) if base_workload is None else np.asarray(base_workload, np.float64)
        assert self.base.shape == (self.W, PERIOD)
        self.kmax = int(math.floor(0.8 * self.W))          # ENV:158 randint(1,80) at W=100
        self.t = 0
        # WORKER:39  initial jittered workload uses the step-0 jitter stream
        self.lw = np.stack([jitter_workload(self.base, _Draws(seed, env0 + e, 0, self.W).jitter())
                            for e in range(self.B)])
        self.R = np.zeros(self.B, np.float64)
        self.C = np.zeros(self.B, np.float64)
        self.arrive = np.zeros(self.B, np.int64)
        self.k_dis = np.zeros(self.B, np.int64)
        self.Pr = np.zeros((self.B, self.W), np.float64)
        self.disabled = np.zeros((self.B, self.W), bool)
        self.n_agents = self.L
        self.observation_space = [[7]] * self.L
        self.share_observation_space = [[2 * self.W + 2 if shannon else self.W + 2]]

    # ---- reset (ENV:157-274, MASTER:46-56, WORKER:114-135) -------------------------------
    def _reset_env(self, e, draw, r):
        W = self.W
        x = draw.reset_scalars(r)
        self.k_dis[e] = 1 + int(px.bounded(x[0], self.kmax))                      # ENV:158
        self.arrive[e] = int(px.bounded(x[1], PERIOD))                            # ENV:160
        self.R[e] = R_MIN + int(px.bounded(x[2], R_HI - R_MIN + 1))               # MASTER:47
        self.C[e] = C_MIN + int(px.bounded(x[3], C_HI - C_MIN + 1))               # MASTER:48
        u, key = draw.worker_block(r)
        self.Pr[e] = 0 + (PR_MAX - 0) * u                                         # ENV:197
        order = np.lexsort((np.arange(W), key))                                   # ENV:199 (choice w/o replacement)
        dis = np.zeros(W, bool)
        dis[order[: self.k_dis[e]]] = True
        self.disabled[e] = dis
        if self.shannon:                                                          # BASE:18-27 (shannon_enable=True)
            pm, pw, d = draw.shannon_block(r)
            up, down = shannon_rates(W, 50 + (60 - 50) * pm, 10 + (20 - 10) * pw, 10 + (100 - 10) * d)
            self.rate_up[e], self.rate_down[e] = up, down

    def observe(self, e, binary=False):
        """ENV:180-271 -> obs (L,7), share_obs row (W+2,), ava (L,2).  ``binary`` (ENV:164-172): R and C as 32-bit strings
        at the head of share_obs (row 62 wider); observation features 0 / 1 are that row's first two entries (ENV:203)."""
        W, a = self.W, int(self.arrive[e])
        Rn = (self.R[e] - R_MIN) / (R_MAX - R_MIN)                                # MASTER:54
        Cn = (self.C[e] - C_MIN) / (C_MAX - C_MIN)
        obs = np.zeros((self.L, 7), np.float64)
        ava = np.ones((self.L, 2), np.int64)
        n_en = W - int(self.k_dis[e])
        d = 0
        prev7 = 0.0
        sums = np.zeros(3)
        sum_pr = 0.0
        for i in range(W):
            lw = self.lw[e, i]
            if self.disabled[e, i]:
                row = [Rn, Cn, 1, 1, 1, 1, prev7]                                 # ENV:207-213
                ava[i, 1] = 0                                                     # WORKER:144-148
                d += 1
            else:
                f7 = (i - d) / (W - int(self.k_dis[e]))                           # ENV:226
                w3 = [lw[(a + s) % PERIOD] for s in range(3)]                     # ENV:222
                row = [Rn, Cn, w3[0], w3[1], w3[2], self.Pr[e, i], f7]
                sums += w3
                sum_pr += self.Pr[e, i]
            prev7 = row[6]
            obs[i] = row
        obs[W] = [Rn, Cn, sums[0] / n_en, sums[1] / n_en, sums[2] / n_en, sum_pr / n_en, 1.1]   # ENV:235-239
        head = [Rn, Cn]
        if binary:                                                                # ENV:164-172
            head = [int(c) for c in "{:032b}".format(int(self.R[e]))] + [int(c) for c in "{:032b}".format(int(self.C[e]))]
            obs[:, 0], obs[:, 1] = head[0], head[1]                               # ENV:203,246 shared_obs[:2]
        if self.shannon:
            sob = np.concatenate([head, self.rate_up[e] / 10 ** 7, self.rate_down[e] / 10 ** 7])   # ENV:248-250
        else:
            sob = np.concatenate([head, self.Pr[e]])                              # ENV:252-253
        return obs, sob, ava

    def _observe_all(self, binary=False):
        o, s, a = zip(*[self.observe(e, binary) for e in range(self.B)])
        sob = np.stack(s)[:, None, :].repeat(self.L, axis=1)                      # ENV:270 replicate over agents
        return np.stack(o), sob, np.stack(a)

    def reset(self, binary=False):
        # a reset of a live env draws a fresh context from the next step's named streams and keeps the workload tables
        # (Env.reset on an existing env, ENV:157-274); only the first reset is "step 0"
        if getattr(self, "_was_reset", False):
            self.t += 1
        self._was_reset = True
        for e in range(self.B):
            self._reset_env(e, _Draws(self.seed, self.env0 + e, self.t, self.W), 0)
        return self._observe_all(binary)

    def set_context(self, R=None, C=None, Pr=None, k_dis=None, disabled=None, arrive=None):
        """Preset contexts (ENV:174-178,191-194 ``preset=True``); arrays indexed by env."""
        for name, val in (("R", R), ("C", C), ("Pr", Pr), ("k_dis", k_dis), ("disabled", disabled), ("arrive", arrive)):
            if val is not None:
                getattr(self, name)[...] = val
        return self._observe_all()

    # ---- step (ENV:57-144) ----------------------------------------------------------------
    def step_env(self, e, action, draw):
        W = self.W
        action = np.asarray(action).reshape(-1)
        if self.fixed:                                                            # ENV:58-62: every available worker, K = 70 %
            sel = np.array([0 if d else 1 for d in self.disabled[e]])
            N = int(sum(sel))
            K = int(np.floor(N * 0.7))
        else:
            ratio = action[-1]
            N = int(sum(action[:-1]))                                             # ENV:71
            K = int(math.ceil(N * ratio))                                         # ENV:72 (fp32 product if fp32 actions)
            sel = action[:-1]
        arrive = int(self.arrive[e])
        R, C = self.R[e], self.C[e]
        R = int(R) if float(R).is_integer() else float(R)
        C = int(C) if float(C).is_integer() else float(C)
        jit = draw.jitter()
        if self.standalone or N == 0:                                             # ENV:81-92
            r = math.ceil(R / 1)
            _, delay, prices = process(self.lw[e, 0], self.Pr[e, 0], r, C, arrive, draw, 0, self.rate_down[e, 0])
            self.lw[e, 0] = jitter_workload(self.base[0], jit[0])
            reward = 1.5 * (-(delay * ALPHA) - (prices[-1] * BETA))
            return reward, delay, prices[-1]
        N = min(max(N, 1), W)                                                     # ENV:96-103
        K = min(max(K, 1), N)
        r = math.ceil(R / K)                                                      # MASTER:39-40
        delays, costs = [], []
        for i in range(W):                                                        # ENV:106-111
            _, dl, pr = process(self.lw[e, i], self.Pr[e, i], r, C, arrive, draw, i, self.rate_down[e, i])
            delays.append(dl)
            costs.append(pr)
        self.lw[e] = jitter_workload(self.base, jit)                              # WORKER:111
        picked = sorted(dl for dl, s in zip(delays, sel) if s != 0)               # ENV:128-129
        final = picked[K - 1]
        end = math.ceil(final)                                                    # ENV:130
        pay = 0.0
        for s, pr in zip(sel, costs):                                             # ENV:131-137
            pay += float(s) * (pr[end - 1] if len(pr) >= end else pr[-1])
        return -(final * ALPHA) - (pay * BETA), final, pay

    def step(self, actions):
        actions = np.asarray(actions).reshape(self.B, self.L)
        self.t += 1
        rew = np.zeros((self.B, self.L, 1), np.float64)
        dones = np.zeros((self.B, self.L), bool)
        infos = []
        for e in range(self.B):
            draw = _Draws(self.seed, self.env0 + e, self.t, self.W)
            reward, delay, pay = self.step_env(e, actions[e], draw)
            self._reset_env(e, draw, 0)                                           # ENV:139
            done = draw.done_coin() < DONE_P                                      # ENV:142
            if done and self.double_reset:                                        # VEC:735-736
                self._reset_env(e, draw, 1)
            rew[e, :, 0] = reward                                                 # BASE:9-13
            dones[e, :] = done
            infos.append([{"delay": delay, "payment": pay}])
        obs, sob, ava = self._observe_all()
        return obs, sob, rew, dones, infos, ava
