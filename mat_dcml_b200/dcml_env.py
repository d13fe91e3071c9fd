"""Vectorised DCML environment behind the reference's vec-env contract
(mat_src/mat/envs/env_wrappers.py:713-744 ShareDummyVecEnv over
DCML_BID_FIRST_MA_ENV_SingleProcess.py:15-355): ``reset() -> (obs, share_obs, ava)``,
``step(actions (N,L,1)) -> (obs, share_obs, rews (N,L,1), dones (N,L), infos, ava)``.

All N envs live on the GPU and advance in one kernel launch (libmatas_b200.so, csrc/env.cu).
``step`` / ``reset`` return host numpy arrays like the reference; ``step_device`` /
``reset_device`` return device tensors and are what a device-resident rollout loop uses
(``share_obs`` is then a stride-0 expand of the one (W+2) row per env).
"""
import ctypes

import numpy as np
import torch

from . import _cabi

PERIOD = 20


class Action_Space:
    """DCML_ENVs/DCML_utils/DCML_ActionSpace.py:1-9 (descriptor the policy dispatches on)."""

    def __init__(self, dim, semi_index=0, extra=False, multi_discrete=False, high=None, low=None, mixed=True,
                 continuous=True):
        self.continuous, self.mixed, self.n, self.extra = continuous, mixed, dim, extra
        self.semi_index, self.multi_discrete, self.high, self.low = semi_index, multi_discrete, high, low


def synthetic_workloads(W, seed=1):
    """``data/workloads.txt``-shaped base table: W rows x 20 slots, multiples of 0.05 in [0,1],
    row means ~0.5, no all-ones row."""
    rng = np.random.RandomState(seed)
    centre = rng.uniform(0.45, 0.6, size=(W, 1))
    raw = np.clip(centre + 0.28 * rng.standard_normal((W, PERIOD)), 0.0, 1.0)
    idx = np.rint(raw / 0.05).astype(np.int64)
    idx[:, 0] = np.minimum(idx[:, 0], 16)
    return (np.arange(21) * 0.05)[idx]


def load_workloads(path, W):
    """Read the reference's ``workloads.txt`` (W consecutive ``np.save`` records, …SingleProcess.py:33-37)."""
    rows = []
    with open(path, "rb") as f:
        for _ in range(W):
            rows.append(np.load(f, allow_pickle=True))
    return np.asarray(rows, np.float64)


class _Infos:
    """Lazy ``[[{"delay":…, "payment":…}], …]`` (dcml_runner.py:63-64 reads info[0]["delay"])."""

    def __init__(self, delay, payment):
        self._d, self._p = delay, payment

    def __len__(self):
        return len(self._d)

    def __getitem__(self, i):
        return [{"delay": float(self._d[i]), "payment": float(self._p[i])}]

    def __iter__(self):
        return (self[i] for i in range(len(self)))


class DcmlVecEnv:
    def __init__(self, n_envs, n_workers=100, base_workload=None, seed=1, device="cuda", double_reset=True,
                 env_offset=0, host_dtype=np.float64, replicate_share_obs=True, shannon=False, fixed=False, standalone=False,
                 multi_agent=True):
        """``host_dtype`` / ``replicate_share_obs`` only shape what ``reset`` / ``step`` hand back on the
        host: the reference returns float64 observations and L materialised copies of the share_obs row
        (…SingleProcess.py:270); float32 and a broadcast view avoid 0.35 GB of host copies per step at
        16384 x 50."""
        self.host_dtype, self.replicate_share_obs = host_dtype, replicate_share_obs
        self.shannon = bool(shannon)      # Shannon link rates on every reset (reference: reset(shannon_enable=True))
        # Env(fixed=True): actions ignored, all available workers, K = floor(0.7 N) (…SingleProcess.py:58-62);
        # standalone: every step is the reference's step(standalone=True) (uncoded on worker 0, reward x 1.5, :81-92)
        self.fixed, self.standalone = bool(fixed), bool(standalone)
        # Env(multi_agent=False) (…SingleProcess.py:42-47, 270-274; DCML_Basic_Env.py:14-17): ONE learner sees the whole env --
        # the host contract hands out obs as one (L*7) vector per env, the share_obs row once, reward (N, 1), dones (N, 1);
        # the simulator itself is unchanged
        self.multi_agent = bool(multi_agent)
        self.num_envs = self.N = int(n_envs)
        self.W, self.L = int(n_workers), int(n_workers) + 1
        self.n_agents = self.L
        self.env_offset = int(env_offset)
        self.seed, self.double_reset = int(seed), bool(double_reset)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _cabi.MatasError("DcmlVecEnv needs a CUDA device (there is no CPU fallback)")
        base = synthetic_workloads(self.W) if base_workload is None else np.asarray(base_workload, np.float64)
        if base.shape != (self.W, PERIOD):
            raise ValueError("base_workload must be (%d, %d)" % (self.W, PERIOD))
        self.base_workload = np.ascontiguousarray(base)
        self.observation_space = [[7]] * self.L if self.multi_agent else [[7 * self.L]]     # …SingleProcess.py:41-44
        self.sob_dim = 2 * self.W + 2 if self.shannon else self.W + 2
        self.share_observation_space = [[self.sob_dim]]               # :45
        self.action_space = ([Action_Space(2, semi_index=-1, extra=True)] if self.multi_agent else   # :46-49
                             [Action_Space(2, semi_index=-1, extra=True, mixed=True, low=0, high=self.W)])
        cfg = _cabi.EnvCfg(self.N, self.W, int(seed) & 0xFFFFFFFFFFFFFFFF, int(env_offset), int(bool(double_reset)), int(self.shannon),
                           int(self.fixed), int(self.standalone))
        h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().dcml_env_create(ctypes.byref(cfg), self.base_workload.ctypes.data_as(ctypes.c_void_p),
                                                    ctypes.byref(h)))
        self._h = h
        f32 = dict(dtype=torch.float32, device=self.device)
        self.obs = torch.empty(self.N, self.L, 7, **f32)
        self.sob = torch.empty(self.N, self.sob_dim, **f32)
        self.ava = torch.empty(self.N, self.L, 2, **f32)
        self.reward = torch.empty(self.N, **f32)
        self.done = torch.empty(self.N, dtype=torch.uint8, device=self.device)
        self.delay = torch.empty(self.N, **f32)
        self.payment = torch.empty(self.N, **f32)
        self._h_act = torch.empty(self.N, self.L, dtype=torch.float32, pin_memory=True)

    def close(self):
        h, self._h = getattr(self, "_h", None), None
        if h is not None and _cabi is not None and _cabi._lib is not None:   # may run at interpreter shutdown
            _cabi._lib.dcml_env_destroy(h)

    __del__ = close

    # ---- device-resident API ----------------------------------------------------------------
    def reset_device(self):
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().dcml_env_reset(self._h, _cabi.ptr(self.obs), _cabi.ptr(self.sob), _cabi.ptr(self.ava),
                                                   _cabi.stream_ptr()))
        return self.obs, self.share_obs_view(), self.ava

    def step_device(self, actions):
        """actions: device fp32 (N, L, 1) or (N, L).  Returns views of the env's own output buffers
        (overwritten by the next call): obs, share_obs, reward (N,), done (N,) uint8, delay, payment, ava."""
        a = actions.to(device=self.device, dtype=torch.float32).reshape(self.N, self.L).contiguous()
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().dcml_env_step(self._h, _cabi.ptr(a), _cabi.ptr(self.obs), _cabi.ptr(self.sob),
                                                  _cabi.ptr(self.ava), _cabi.ptr(self.reward), _cabi.ptr(self.done),
                                                  _cabi.ptr(self.delay), _cabi.ptr(self.payment), _cabi.stream_ptr()))
        return self.obs, self.share_obs_view(), self.reward, self.done, self.delay, self.payment, self.ava

    def share_obs_view(self):
        return self.sob.unsqueeze(1).expand(self.N, self.L, self.sob_dim)

    def set_context(self, R=None, C=None, worker_pr=None, n_disabled=None):
        """Preset contexts (Env(preset=True) / modify_preset); arrays per env, None keeps the current value."""
        def dev(x, dt, shape):
            if x is None:
                return None
            return torch.as_tensor(np.broadcast_to(np.asarray(x), shape).copy()).to(device=self.device, dtype=dt).contiguous()
        R = dev(R, torch.float64, (self.N,))
        C = dev(C, torch.float64, (self.N,))
        pr = dev(worker_pr, torch.float64, (self.N, self.W))
        if n_disabled is not None:
            k = np.asarray(n_disabled)
            if k.min() < 0 or k.max() >= self.W:      # W - k enabled workers divide the 7th feature and the master row means
                raise ValueError("n_disabled must be in [0, %d)" % self.W)
        nd = dev(n_disabled, torch.int32, (self.N,))
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().dcml_env_set_context(self._h, _cabi.ptr(R), _cabi.ptr(C), _cabi.ptr(pr), _cabi.ptr(nd),
                                                         _cabi.ptr(self.obs), _cabi.ptr(self.sob), _cabi.ptr(self.ava),
                                                         _cabi.stream_ptr()))
        return self.obs, self.share_obs_view(), self.ava

    # ---- the reference's host-side contract -------------------------------------------------
    def _final(self, name, s0, s1):
        """Rows [s0, s1) of output `name` as the DEVICE tensor the host contract hands out: dtype conversion (the reference
        returns float64 observations / rewards and int64 flags) and the replication of the per-env share_obs row, reward and
        done flag over the L agents (…SingleProcess.py:270, DCML_Basic_Env.py:9-13) are done on the GPU, so that the host
        side is one pinned copy per array and no numpy pass over hundreds of megabytes."""
        f64 = self.host_dtype == np.float64
        dt = torch.float64 if f64 else torch.float32
        ma, rep, n = self.multi_agent, self.replicate_share_obs and self.multi_agent, s1 - s0
        src = getattr(self, name)[s0:s1]
        if name == "obs":
            x = src.to(dt)
            return x if ma else x.reshape(n, self.L * 7)
        if name in ("sob", "sob_bits"):
            x = src.to(dt)
            return x.unsqueeze(1).expand(n, self.L, x.shape[-1]).contiguous() if rep else x
        if name == "ava":
            return src.to(torch.int64 if f64 else torch.float32)
        if name == "reward":
            x = src.to(dt)
            return x[:, None, None].expand(n, self.L, 1).contiguous() if rep else x
        if name == "done":
            x = src.bool()
            return x[:, None].expand(n, self.L).contiguous() if rep else x
        return src

    def _fetch(self, names):
        """Device -> host copies into FRESH pinned tensors (torch's caching host allocator recycles the blocks), so the
        numpy arrays handed out are zero-copy views the caller owns, as with the reference's freshly allocated arrays."""
        host = {}
        for k in names:
            x = self._final(k, 0, self.N)
            host[k] = torch.empty(x.shape, dtype=x.dtype, pin_memory=True).copy_(x, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return host

    def _host_obs(self, host, sob_key="sob"):
        obs, sob, ava = host["obs"].numpy(), host[sob_key].numpy(), host["ava"].numpy()
        if self.multi_agent and not self.replicate_share_obs:
            sob = np.broadcast_to(sob[:, None, :], (self.N, self.L, sob.shape[-1]))
        return obs, sob, ava

    HOST_SLICES = 4            # a large batch is stepped in this many env slices when the results go to the host
    HOST_SLICE_MIN = 2048

    def _step_sliced(self, a_dev, names):
        """One step issued as env slices (dcml_env_step_slice): the device-to-host copies of a slice's output arrays run on
        a copy stream under the next slice's kernel.  Same kernel, same named draws, same results as step_device."""
        dev = self.device
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
        a_dev = a_dev.to(dtype=torch.float32).reshape(self.N, self.L).contiguous()
        n = self.HOST_SLICES
        bounds = [self.N * k // n for k in range(n + 1)]
        lib = _cabi.lib()
        host = {}
        cur, cs = torch.cuda.current_stream(dev), self._copy_stream
        with torch.cuda.device(dev):
            for k in range(n):
                s0, s1 = bounds[k], bounds[k + 1]
                _cabi.check(lib.dcml_env_step_slice(self._h, _cabi.ptr(a_dev), s0, s1 - s0, int(k == 0), _cabi.ptr(self.obs), _cabi.ptr(self.sob),
                                                    _cabi.ptr(self.ava), _cabi.ptr(self.reward), _cabi.ptr(self.done), _cabi.ptr(self.delay),
                                                    _cabi.ptr(self.payment), _cabi.stream_ptr()))
                parts = {name: self._final(name, s0, s1) for name in names}
                ev = cur.record_event()
                with torch.cuda.stream(cs):
                    cs.wait_event(ev)
                    for name, x in parts.items():
                        if name not in host:
                            host[name] = torch.empty((self.N,) + tuple(x.shape[1:]), dtype=x.dtype, pin_memory=True)
                        host[name][s0:s1].copy_(x, non_blocking=True)
                        x.record_stream(cs)
        cur.wait_stream(cs)            # later kernels overwrite the output buffers: order them behind the copies
        cs.synchronize()
        return host

    def reset(self, binary=False):
        """Env.reset through the vec-env.  ``binary=True`` is the reference's ``reset(binary=True)``
        (…SingleProcess.py:164-172): R and C enter share_obs as two 32-bit strings, which makes its rows 62 wider, and the
        first two features of every observation row are that row's first two entries.  Only this observation is binary: a
        step's own reset is not (:139)."""
        if not binary:
            self.reset_device()
            return self._host_obs(self._fetch(("obs", "sob", "ava")))
        self.sob_bits = torch.empty(self.N, self.sob_dim + 62, dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().dcml_env_reset_binary(self._h, _cabi.ptr(self.obs), _cabi.ptr(self.sob_bits), _cabi.ptr(self.ava),
                                                          _cabi.stream_ptr()))
        return self._host_obs(self._fetch(("obs", "sob_bits", "ava")), "sob_bits")

    def step(self, actions):
        if not torch.is_tensor(actions):
            actions = torch.from_numpy(np.ascontiguousarray(actions, dtype=np.float32))
        actions = actions.reshape(self.N, self.L)
        if actions.device.type == "cpu" and not (actions.dtype == torch.float32 and actions.is_contiguous() and actions.is_pinned()):
            self._h_act.copy_(actions)             # pageable / strided / non-fp32 input: stage through pinned memory
            actions = self._h_act
        names = ("obs", "sob", "ava", "reward", "done", "delay", "payment")
        if self.N >= self.HOST_SLICE_MIN:
            host = self._step_sliced(actions.to(self.device, non_blocking=True), names)
        else:
            self.step_device(actions.to(self.device, non_blocking=True))
            host = self._fetch(names)
        obs, sob, ava = self._host_obs(host)
        rew, dones = host["reward"].numpy(), host["done"].numpy()
        infos = _Infos(host["delay"].numpy(), host["payment"].numpy())
        if not self.multi_agent:                                      # DCML_Basic_Env.py:14-17: one reward / done per env
            return obs, sob, rew[:, None], dones[:, None], infos, ava
        if not self.replicate_share_obs:
            rew, dones = np.broadcast_to(rew[:, None, None], (self.N, self.L, 1)), np.broadcast_to(dones[:, None], (self.N, self.L))
        return obs, sob, rew, dones, infos, ava
