"""Device-resident rollout store with the reference's ``SharedReplayBuffer`` interface
(mat_src/mat/utils/shared_buffer.py:22-314): same array names and (T+1 | T, N, L, dim) shapes,
``insert`` / ``compute_returns`` / ``after_update`` / ``feed_forward_generator_transformer``.

Host arrays handed to ``insert`` (the reference's numpy contract) are uploaded on a side stream when they
live in pinned memory, so the copies overlap the next step's kernels; every read of a buffer array first makes
the current stream wait for pending uploads (the arrays below are properties for that reason).

Differences, all layout-only: arrays are fp32 torch tensors on the GPU instead of numpy;
``share_obs`` keeps ONE (W+2) row per (t, env) and is exposed as a stride-0 expand
(the model ignores it, ma_transformer.py:144-149); the dead ``rnn_states*`` arrays are
(T+1, N, L, 1, 1) zeros.  ``compute_returns`` is the GAE kernel (csrc/gae.cu).
"""
import ctypes
import warnings

import numpy as np
import torch

from . import _cabi


def _shape(space):
    return space.shape if space.__class__.__name__ == "Box" else space


def _as_tensor(x):
    """torch view of a host array without the "not writable" warning: env outputs may be read-only broadcast views (the
    share_obs row / reward replicated over the agents), and the buffer only ever reads them."""
    if torch.is_tensor(x):
        return x
    if isinstance(x, np.ndarray) and not x.flags.writeable:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore", UserWarning)
            return torch.as_tensor(x)
    return torch.as_tensor(x)


def _guarded(name):
    key = "_t_" + name
    def get(self):
        self._join()
        return self.__dict__[key]
    def put(self, value):
        self.__dict__[key] = value
    return property(get, put)


class SharedReplayBuffer:
    # arrays that insert() may fill from the side stream
    _share_obs = _guarded("_share_obs")
    obs = _guarded("obs")
    actions = _guarded("actions")
    action_log_probs = _guarded("action_log_probs")
    value_preds = _guarded("value_preds")
    available_actions = _guarded("available_actions")
    rewards = _guarded("rewards")
    masks = _guarded("masks")
    bad_masks = _guarded("bad_masks")
    active_masks = _guarded("active_masks")

    def __init__(self, args, num_agents, obs_space, cent_obs_space, act_space, env_name, device="cuda"):
        self.episode_length, self.n_rollout_threads = args.episode_length, args.n_rollout_threads
        self.gamma, self.gae_lambda = args.gamma, args.gae_lambda
        self._use_gae, self._use_popart, self._use_valuenorm = args.use_gae, args.use_popart, args.use_valuenorm
        self.num_agents, self.env_name = num_agents, env_name
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _cabi.MatasError("SharedReplayBuffer needs a CUDA device (there is no CPU fallback)")
        # use_gae is stored like the reference does and, like there, never read: compute_returns only has GAE branches
        # (shared_buffer.py:207-238)
        self._copy_stream = torch.cuda.Stream(device=self.device)
        self._dirty = False
        self._inflight = []            # (event, source tensors) of uploads that may still be reading pinned memory
        self._held = []
        T, N, L = self.episode_length, self.n_rollout_threads, num_agents
        obs_dim = _shape(obs_space)[0] if not isinstance(_shape(obs_space)[-1], list) else _shape(obs_space)[0][0]
        sob_dim = _shape(cent_obs_space)[0] if not isinstance(_shape(cent_obs_space)[-1], list) else _shape(cent_obs_space)[0][0]
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=self.device)
        o = lambda *s: torch.ones(*s, dtype=torch.float32, device=self.device)
        self._share_obs = z(T + 1, N, sob_dim)
        self.obs = z(T + 1, N, L, obs_dim)
        self.rnn_states = z(T + 1, N, L, 1, 1)
        self.rnn_states_critic = self.rnn_states
        self.value_preds = z(T + 1, N, L, 1)
        self.returns = z(T + 1, N, L, 1)
        self.advantages = z(T, N, L, 1)
        self.available_actions = o(T + 1, N, L, act_space.n)
        self.actions = z(T, N, L, 1)
        self.action_log_probs = z(T, N, L, 1)
        self.rewards = z(T, N, L, 1)
        self.masks = o(T + 1, N, L, 1)
        self.bad_masks = o(T + 1, N, L, 1)
        self.active_masks = o(T + 1, N, L, 1)
        self.step = 0

    @property
    def share_obs(self):
        T1, N, D = self._share_obs.shape
        return self._share_obs.unsqueeze(2).expand(T1, N, self.num_agents, D)

    def _join(self):
        """Order the current stream after the uploads issued so far."""
        if self.__dict__.get("_dirty"):
            torch.cuda.current_stream(self.device).wait_stream(self._copy_stream)
            self._dirty = False

    def _put(self, name, index, src):
        """array[index] = src.  Pinned fp32 host sources are uploaded on the side stream."""
        dst = self.__dict__["_t_" + name][index]
        src = _as_tensor(src)
        if (src.device.type == "cpu" and src.dtype == torch.float32 and src.is_contiguous() and src.numel() == dst.numel()
                and src.is_pinned()):
            cs = self._copy_stream
            cs.wait_stream(torch.cuda.current_stream(self.device))      # kernels queued earlier may still read dst
            with torch.cuda.stream(cs):
                dst.copy_(src.reshape(dst.shape), non_blocking=True)
            self._held.append(src)
            self._dirty = True
        else:
            self._join()
            dst.copy_(src.to(device=self.device, dtype=torch.float32).reshape(dst.shape), non_blocking=True)

    def insert(self, share_obs, obs, rnn_states_actor, rnn_states_critic, actions, action_log_probs, value_preds,
               rewards, masks, bad_masks=None, active_masks=None, available_actions=None):
        """shared_buffer.py:118-151.  ``share_obs`` may be (N, L, D) (the reference's replicated form;
        row 0 is kept) or (N, D); ``rewards`` / ``masks`` may be per env (N,) or (N, L, 1)."""
        s = self.step
        self._inflight = [(e, h) for e, h in self._inflight if not e.query()]      # drop finished uploads
        self._held = []
        so = _as_tensor(share_obs)
        self._put("_share_obs", s + 1, so[:, 0, :] if so.dim() == 3 else so)
        self._put("obs", s + 1, obs)
        self._put("actions", s, actions)
        self._put("action_log_probs", s, action_log_probs)
        self._put("value_preds", s, value_preds)
        raw = self.__dict__                        # unguarded tensors: _bcast orders its own writes
        self._bcast(raw["_t_rewards"][s], rewards)
        self._bcast(raw["_t_masks"][s + 1], masks)
        if bad_masks is not None:
            self._bcast(raw["_t_bad_masks"][s + 1], bad_masks)
        if active_masks is not None:
            self._bcast(raw["_t_active_masks"][s + 1], active_masks)
        if available_actions is not None:
            self._put("available_actions", s + 1, available_actions)
        if self._held:
            ev = torch.cuda.Event()
            ev.record(self._copy_stream)
            self._inflight.append((ev, self._held))
        self._held = []
        self.step = (self.step + 1) % self.episode_length

    def _bcast(self, dst, src):
        """dst (N, L, 1) = src, given per env (N,) or per agent (N, L, 1).  Host sources are staged through a small pinned
        tensor and uploaded on the side stream (a pageable source would make the copy synchronous with the host)."""
        src = _as_tensor(src)
        if src.device.type == "cpu":
            pin = torch.empty(src.shape, dtype=torch.float32, pin_memory=True)
            pin.copy_(src)
            cs = self._copy_stream
            cs.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(cs):
                d = pin.to(self.device, non_blocking=True)
                d = d.reshape(-1, 1, 1).expand_as(dst) if d.numel() == dst.shape[0] else d.reshape(dst.shape)
                dst.copy_(d)
            self._held.append(pin)
            self._dirty = True
            return
        src = src.to(device=self.device, dtype=torch.float32)
        if src.numel() == dst.shape[0]:
            src = src.reshape(-1, 1, 1).expand_as(dst)
        dst.copy_(src.reshape(dst.shape) if src.numel() == dst.numel() else src)

    def after_update(self):
        """shared_buffer.py:188-198"""
        for a in (self._share_obs, self.obs, self.masks, self.bad_masks, self.active_masks, self.available_actions):
            a[0].copy_(a[-1])

    def compute_returns(self, next_value, value_normalizer=None):
        """shared_buffer.py:207-238 (GAE branch) on the device, ValueNorm denormalisation fused."""
        self._held = []
        self._put("value_preds", -1, next_value)
        if self._held:
            ev = torch.cuda.Event()
            ev.record(self._copy_stream)
            self._inflight.append((ev, self._held))
            self._held = []
        use_vn = (self._use_popart or self._use_valuenorm) and value_normalizer is not None
        T, N, L = self.episode_length, self.n_rollout_threads, self.num_agents
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().matas_gae(_cabi.ptr(self.rewards), _cabi.ptr(self.value_preds), _cabi.ptr(self.masks),
                                              _cabi.ptr(value_normalizer.stats) if use_vn else None,
                                              float(self.gamma), float(self.gae_lambda), T, N * L,
                                              _cabi.ptr(self.advantages), _cabi.ptr(self.returns), _cabi.stream_ptr()))

    def feed_forward_generator_transformer(self, advantages, num_mini_batch=None, mini_batch_size=None):
        """shared_buffer.py:240-314: rows are (t, env) pairs carrying all L agents; one global
        permutation split into contiguous minibatches; yields the reference's 12-tuple (device tensors)."""
        T, N, L = self.episode_length, self.n_rollout_threads, self.num_agents
        batch_size = N * T
        if mini_batch_size is None:
            assert batch_size >= num_mini_batch, "PPO requires N*T >= num_mini_batch"
            mini_batch_size = batch_size // num_mini_batch
        rand = torch.randperm(batch_size, device=self.device)
        flat = lambda x: x.reshape(batch_size, -1)
        adv = None if advantages is None else torch.as_tensor(advantages).to(device=self.device, dtype=torch.float32).contiguous()
        # (source viewed as [batch][row], floats per row, copies per row): share_obs is kept as one row per sample and
        # replicated over the agents on the way out, everything else carries its L agents in the row
        srcs = [(flat(self._share_obs[:-1]), L), (flat(self.obs[:-1]), 1), (flat(self.actions), 1), (flat(self.value_preds[:-1]), 1),
                (flat(self.returns[:-1]), 1), (flat(self.masks[:-1]), 1), (flat(self.active_masks[:-1]), 1),
                (flat(self.action_log_probs), 1), (flat(self.available_actions[:-1]), 1)]
        if adv is not None:
            srcs.append((flat(adv), 1))
        n = len(srcs)
        c_src = (ctypes.c_void_p * n)(*[s.data_ptr() for s, _ in srcs])
        c_w = (ctypes.c_int * n)(*[s.shape[1] for s, _ in srcs])
        c_rep = (ctypes.c_int * n)(*[r for _, r in srcs])
        for i in range(num_mini_batch):
            idx = rand[i * mini_batch_size:(i + 1) * mini_batch_size].contiguous()
            M = idx.numel()
            # (M * L, dim) outputs, as the reference's `x[indices].reshape(-1, *x.shape[2:])`
            outs = [torch.empty(M * L, s.shape[1] * r // L, dtype=torch.float32, device=self.device) for s, r in srcs]
            c_dst = (ctypes.c_void_p * n)(*[o.data_ptr() for o in outs])
            with torch.cuda.device(self.device):
                _cabi.check(_cabi.lib().matas_gather_rows(n, c_src, c_dst, c_w, c_rep, _cabi.ptr(idx), M, _cabi.stream_ptr()))
            so, ob, act, vp, ret, msk, am, lp, ava = outs[:9]
            yield (so, ob, None, None, act, vp, ret, msk, am, lp, outs[9] if adv is not None else None, ava)
