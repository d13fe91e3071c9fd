"""Drop-in for the reference's ``MultiAgentTransformer``
(mat_src/mat/algorithms/mat/algorithm/ma_transformer.py:233-339) whose arithmetic runs in the
hand-written sm_100a kernels of libmatas_b200.so.

The ``nn.Module`` tree below exists for ONE reason: to own parameters under exactly the
reference's names so that ``state_dict()`` / ``load_state_dict()`` / ``torch.save`` round-trip
the reference's checkpoints (120 keys at the DCML shape, including the six ``attn*.mask``
buffers and the never-used ``encoder.state_encoder.*`` / ``decoder.obs_encoder.*``,
SURVEY.md App. D #8) and so that optimizers and ``clip_grad_norm_`` see ordinary parameters.
No module here has a torch ``forward``: there is no eager fallback.  All parameters are views
into one flat fp32 buffer in state_dict order, which is what the C ABI consumes
(``matas_policy_set_weights``) and what a gradient all-reduce operates on.
"""
import ctypes
import math

import numpy as np
import torch
import torch.nn as nn

from . import _cabi


def _init(m, gain=0.01, activate=False):
    # ma_transformer.py:18-21: orthogonal weights (gain sqrt(2) before an activation), zero bias
    if activate:
        gain = nn.init.calculate_gain("relu")
    nn.init.orthogonal_(m.weight.data, gain=gain)
    if m.bias is not None:
        nn.init.constant_(m.bias.data, 0)
    return m


class _Holder(nn.Module):
    def forward(self, *a, **k):
        raise RuntimeError("mat_dcml_b200 modules only hold parameters; the forward pass lives in "
                           "libmatas_b200.so (call the MultiAgentTransformer / TransformerPolicy methods)")


class SelfAttention(_Holder):
    def __init__(self, n_embd, n_head, n_agent, masked=False):
        super().__init__()
        self.masked, self.n_head = masked, n_head
        self.key = _init(nn.Linear(n_embd, n_embd))
        self.query = _init(nn.Linear(n_embd, n_embd))
        self.value = _init(nn.Linear(n_embd, n_embd))
        self.proj = _init(nn.Linear(n_embd, n_embd))
        self.register_buffer("mask", torch.tril(torch.ones(n_agent + 1, n_agent + 1)).view(1, 1, n_agent + 1, n_agent + 1))


def _mlp(n_embd):
    return nn.Sequential(_init(nn.Linear(n_embd, n_embd), activate=True), nn.GELU(), _init(nn.Linear(n_embd, n_embd)))


class EncodeBlock(_Holder):
    def __init__(self, n_embd, n_head, n_agent):
        super().__init__()
        self.ln1 = nn.LayerNorm(n_embd)
        self.ln2 = nn.LayerNorm(n_embd)
        self.attn = SelfAttention(n_embd, n_head, n_agent, masked=False)
        self.mlp = _mlp(n_embd)


class DecodeBlock(_Holder):
    def __init__(self, n_embd, n_head, n_agent):
        super().__init__()
        self.ln1 = nn.LayerNorm(n_embd)
        self.ln2 = nn.LayerNorm(n_embd)
        self.ln3 = nn.LayerNorm(n_embd)
        self.attn1 = SelfAttention(n_embd, n_head, n_agent, masked=True)
        self.attn2 = SelfAttention(n_embd, n_head, n_agent, masked=True)
        self.mlp = _mlp(n_embd)


def _head(n_embd, out):
    return nn.Sequential(_init(nn.Linear(n_embd, n_embd), activate=True), nn.GELU(), nn.LayerNorm(n_embd),
                         _init(nn.Linear(n_embd, out)))


def _embed(in_dim, n_embd):
    return nn.Sequential(nn.LayerNorm(in_dim), _init(nn.Linear(in_dim, n_embd), activate=True), nn.GELU())


class Encoder(_Holder):
    def __init__(self, state_dim, obs_dim, n_block, n_embd, n_head, n_agent, encode_state):
        super().__init__()
        self.state_dim, self.obs_dim, self.n_embd, self.n_agent, self.encode_state = state_dim, obs_dim, n_embd, n_agent, encode_state
        self.state_encoder = _embed(state_dim, n_embd)
        self.obs_encoder = _embed(obs_dim, n_embd)
        self.ln = nn.LayerNorm(n_embd)
        self.blocks = nn.Sequential(*[EncodeBlock(n_embd, n_head, n_agent) for _ in range(n_block)])
        self.head = _head(n_embd, 1)


class Decoder(_Holder):
    def __init__(self, obs_dim, action_dim, n_block, n_embd, n_head, n_agent, action_type="Discrete"):
        super().__init__()
        self.action_dim, self.n_embd, self.action_type = action_dim, n_embd, action_type
        self.dec_actor = self.share_actor = False
        if action_type != "Discrete":
            self.log_std = nn.Parameter(torch.ones(action_dim))
        self.action_encoder = nn.Sequential(_init(nn.Linear(action_dim + 1, n_embd, bias=False), activate=True), nn.GELU())
        self.obs_encoder = _embed(obs_dim, n_embd)
        self.ln = nn.LayerNorm(n_embd)
        self.blocks = nn.Sequential(*[DecodeBlock(n_embd, n_head, n_agent) for _ in range(n_block)])
        self.head = _head(n_embd, action_dim)
        self._owner = None      # set by MultiAgentTransformer; lets transformer_act reach the handle

    def zero_std(self, device=None):
        if self.action_type != "Discrete":
            self.log_std.data.zero_()


def _check(x):
    return torch.from_numpy(x) if isinstance(x, np.ndarray) else x


class MultiAgentTransformer(nn.Module):
    """Same constructor and methods as the reference class; see module docstring."""

    def __init__(self, state_dim, obs_dim, action_dim, n_agent, n_block, n_embd, n_head, encode_state=False,
                 device=torch.device("cpu"), action_type="Discrete", dec_actor=False, share_actor=False, semi_index=-1):
        super().__init__()
        if encode_state or dec_actor or share_actor:
            raise NotImplementedError("encode_state / dec_actor / share_actor are outside the MAT-AS hot path "
                                      "(no DCML config enables them, mat_src/mat/config.py:304-310)")
        if action_type not in ("Discrete", "Semi_Discrete"):
            raise NotImplementedError("action_type %r: only the Discrete and Semi_Discrete act paths are built "
                                      "(ma_transformer.py:321-328 are unreachable from DCML)" % (action_type,))
        if action_type == "Semi_Discrete" and semi_index != -1:
            raise NotImplementedError("semi_index must be -1 (one continuous agent, DCML_Config.py:3)")
        self.n_agent, self.action_dim, self.obs_dim, self.state_dim = n_agent, action_dim, obs_dim, state_dim
        self.n_block, self.n_embd, self.n_head = n_block, n_embd, n_head
        self.action_type, self.semi_index = action_type, semi_index
        self.device = torch.device(device)
        self.tpdv = dict(dtype=torch.float32, device=self.device)
        self.encoder = Encoder(state_dim, obs_dim, n_block, n_embd, n_head, n_agent, encode_state)
        self.decoder = Decoder(obs_dim, action_dim, n_block, n_embd, n_head, n_agent, action_type)
        self.decoder._owner = [self]          # list: keep nn.Module from registering a cycle
        self.to(self.device)
        self._handle = None
        self._flat = None
        self._flat_grad = None
        self._flatten()
        # named Philox draws for stochastic rollouts (DESIGN.md "Random streams")
        self.rollout_seed = int(torch.initial_seed()) & 0xFFFFFFFFFFFFFFFF
        self.rollout_step = 0
        self.env_offset = 0
        self.output_device = torch.device("cpu")   # reference returns CPU actions / log-probs (transformer_act.py:34-35)

    # ---- flat parameter storage -------------------------------------------------------------
    def _flatten(self):
        params = list(self.parameters())
        n = sum(p.numel() for p in params)
        flat = torch.empty(n, dtype=torch.float32, device=self.device)
        off = 0
        for p in params:
            k = p.numel()
            flat[off:off + k].copy_(p.data.reshape(-1))
            p.data = flat[off:off + k].view(p.shape)
            off += k
        self._flat = flat
        self._plist = params
        self._packed_key = None

    def flat_parameters(self):
        return self._flat

    def flat_grad(self, zero=False):
        """One flat fp32 gradient buffer; every parameter's .grad is a view into it."""
        if self._flat_grad is None:
            # four spare slots behind the gradient carry the loss denominators / numerators of a sharded minibatch through
            # the gradient all-reduce (mat_trainer.ppo_update); flat_grad() is the gradient itself
            self._flat_grad_comm = torch.zeros(self._flat.numel() + 4, dtype=torch.float32, device=self._flat.device)
            self._flat_grad = self._flat_grad_comm[:self._flat.numel()]
            off = 0
            for p in self.parameters():
                k = p.numel()
                p.grad = self._flat_grad[off:off + k].view(p.shape)
                off += k
        elif zero:
            self._flat_grad.zero_()
        return self._flat_grad

    def flat_grad_comm(self):
        """The gradient buffer with its four spare slots (see flat_grad)."""
        self.flat_grad()
        return self._flat_grad_comm

    def zero_std(self):
        self.decoder.zero_std()

    def _anchor(self):
        """A requires-grad scalar that ties the CUDA backward into the autograd graph."""
        a = getattr(self, "_anchor_t", None)
        if a is None or a.device != self.device:
            a = self._anchor_t = torch.zeros((), device=self.device, requires_grad=True)
        return a

    # ---- handle -----------------------------------------------------------------------------
    def _ctx(self):
        """Device guard for kernel calls; raises (never falls back) when the model is not on a GPU."""
        if self.device.type != "cuda":
            raise _cabi.MatasError("MultiAgentTransformer is on %s: the MAT-AS kernels need a CUDA device "
                                   "(there is no CPU fallback)" % self.device)
        return torch.cuda.device(self.device)

    def _h(self):
        self._ctx()
        if self._handle is None:
            cfg = _cabi.ModelCfg(self.n_agent, self.obs_dim, self.state_dim, self.action_dim, self.n_block, self.n_embd,
                                 self.n_head, _cabi.MODE_SEMI if self.action_type == "Semi_Discrete" else _cabi.MODE_DISCRETE)
            h = ctypes.c_void_p()
            with self._ctx():
                _cabi.check(_cabi.lib().matas_policy_create(ctypes.byref(cfg), ctypes.byref(h)))
                n = _cabi.lib().matas_policy_param_count(h)
            if n != self._flat.numel():
                raise _cabi.MatasError("parameter count mismatch: module %d vs library %d" % (self._flat.numel(), n))
            self._handle = h
        return self._handle

    def __del__(self):
        h, self._handle = getattr(self, "_handle", None), None
        if h is not None:
            try:
                _cabi.lib().matas_policy_destroy(h)
            except Exception:
                pass

    def sync_weights(self):
        """Repack the flat parameters for the kernels (three small launches).  Called before every pass; skipped
        when neither the flat tensor nor any parameter view has been written since the last repack (torch's version
        counters: optimizer steps, ``load_state_dict`` and any other in-place torch update are seen; code that
        writes the memory behind torch's back, e.g. a collective on ``flat_parameters()``, must call
        ``invalidate_weights``)."""
        h = self._h()
        # parameters are views created with ``p.data = flat[...]``: each keeps its own version counter
        key = (self._flat._version, self._flat.data_ptr(), tuple(p._version for p in self._plist))
        if key != getattr(self, "_packed_key", None):
            _cabi.check(_cabi.lib().matas_policy_set_weights(h, _cabi.ptr(self._flat), _cabi.stream_ptr()))
            self._packed_key = key
        return h

    def invalidate_weights(self):
        self._packed_key = None

    def _dev(self, x, shape=None):
        x = _check(x)
        if not torch.is_tensor(x):
            x = torch.as_tensor(np.asarray(x))
        x = x.to(device=self.device, dtype=torch.float32, non_blocking=True)
        if shape is not None:
            x = x.reshape(shape)
        return x.contiguous()

    def _out(self, *xs):
        """Results on ``output_device``.  Host outputs go through fresh pinned tensors (recycled by torch's
        caching host allocator) with one stream synchronisation for all of them."""
        if self.output_device.type == "cuda":
            return xs if len(xs) > 1 else xs[0]
        hs = tuple(torch.empty(x.shape, dtype=x.dtype, pin_memory=True).copy_(x, non_blocking=True) for x in xs)
        torch.cuda.current_stream(self.device).synchronize()
        return hs if len(hs) > 1 else hs[0]

    def to_host(self, *xs):
        """Device tensors -> pinned host tensors with ONE stream synchronisation (what a host-side caller that wants numpy
        values, like the reference's runner with ``_t2n``, should use instead of one ``.cpu()`` per tensor)."""
        hs = tuple(torch.empty(x.shape, dtype=x.dtype, pin_memory=True).copy_(x, non_blocking=True) for x in xs)
        torch.cuda.current_stream(self.device).synchronize()
        return hs if len(hs) > 1 else hs[0]

    # ---- the reference's methods ------------------------------------------------------------
    def get_values(self, state, obs, available_actions=None):
        """ma_transformer.py:331-339 -> v_loc (B, L, 1) on the model device."""
        obs = self._dev(obs, (-1, self.n_agent, self.obs_dim))
        B = obs.shape[0]
        values = torch.empty(B, self.n_agent, 1, **self.tpdv)
        with self._ctx():
            h = self.sync_weights()
            _cabi.check(_cabi.lib().matas_policy_get_values(h, _cabi.ptr(obs), B, _cabi.ptr(values), _cabi.stream_ptr()))
        return values

    def encode(self, obs):
        """Encoder.forward (ma_transformer.py:141-154) -> (v_loc (B,L,1), obs_rep (B,L,64))."""
        obs = self._dev(obs, (-1, self.n_agent, self.obs_dim))
        B = obs.shape[0]
        values = torch.empty(B, self.n_agent, 1, **self.tpdv)
        rep = torch.empty(B, self.n_agent, self.n_embd, **self.tpdv)
        with self._ctx():
            h = self.sync_weights()
            _cabi.check(_cabi.lib().matas_policy_encode(h, _cabi.ptr(obs), B, _cabi.ptr(values), _cabi.ptr(rep), _cabi.stream_ptr()))
        return values, rep

    HOST_SLICES = 4            # host observations of large batches are uploaded in this many env slices (see _get_actions_host)
    HOST_SLICE_MIN = 2048      # envs below which one copy is as good

    def _get_actions_host(self, obs, ava, deterministic, stride):
        """get_actions for HOST inputs of a large batch (the reference's contract, ma_transformer.py:304-307): the
        observations go up in env slices on a copy stream and the encoder of slice k runs while slice k + 1 is still on
        the bus; the availability flags, which only the decode reads, follow behind the last slice.  Same kernels and same
        results as the one-piece path (matas_policy_encode_slice + matas_policy_decode)."""
        B, L = obs.shape[0], self.n_agent
        dev = self.device
        cur = torch.cuda.current_stream(dev)
        cs = getattr(self, "_copy_stream", None)
        if cs is None:
            cs = self._copy_stream = torch.cuda.Stream(device=dev)
        obs_d = torch.empty(B, L, self.obs_dim, **self.tpdv)
        ava_d = torch.empty(B, L, self.action_dim, **self.tpdv)
        values = torch.empty(B, L, 1, **self.tpdv)
        actions = torch.empty(B, L, 1, **self.tpdv)
        logp = torch.empty(B, L, 1, **self.tpdv)
        obs_d.record_stream(cs); ava_d.record_stream(cs)
        cs.wait_stream(cur)
        step = self.rollout_step
        self.rollout_step = (self.rollout_step + 1) & 0xFFFFFFFF
        n = self.HOST_SLICES
        bounds = [B * k // n for k in range(n + 1)]
        lib = _cabi.lib()
        with self._ctx():
            h = self.sync_weights()
            for k in range(n):
                s0, s1 = bounds[k], bounds[k + 1]
                with torch.cuda.stream(cs):
                    self._upload(obs_d[s0:s1], obs[s0:s1])
                    ev = cs.record_event()
                cur.wait_event(ev)
                _cabi.check(lib.matas_policy_encode_slice(h, _cabi.ptr(obs_d[s0:s1]), s0, s1 - s0, B, _cabi.ptr(values[s0:s1]),
                                                          _cabi.stream_ptr()))
            with torch.cuda.stream(cs):
                self._upload(ava_d, ava)
                ev = cs.record_event()
            cur.wait_event(ev)
            _cabi.check(lib.matas_policy_decode(h, _cabi.ptr(ava_d), B, int(bool(deterministic)), int(stride), self.rollout_seed, step,
                                                self.env_offset, None, None, _cabi.ptr(actions), _cabi.ptr(logp), _cabi.stream_ptr()))
        return actions, logp, values

    @staticmethod
    def _upload(dst, src):
        """dst (device fp32) = src (host) on the current stream.  The env's own host contract is float64 observations and
        int64 flags: those go up as they are and are converted on the device (a dtype-changing copy_ would convert on the
        host first)."""
        if src.dtype == dst.dtype:
            dst.copy_(src, non_blocking=True)
        else:
            stage = torch.empty(src.shape, dtype=src.dtype, device=dst.device)
            stage.copy_(src, non_blocking=True)
            dst.copy_(stage)

    @staticmethod
    def _host_f32(x, shape):
        """x as a contiguous CPU tensor of `shape` (fp32, or the float64 / int64 of the reference's env) without a copy, or None."""
        x = _check(x)
        if not torch.is_tensor(x) or x.device.type != "cpu" or x.dtype not in (torch.float32, torch.float64, torch.int64):
            return None
        x = x.reshape(shape)
        return x if x.is_contiguous() else None

    def get_actions(self, state, obs, available_actions=None, deterministic=False, stride=2, u=None, z=None):
        """ma_transformer.py:298-329 -> (actions (B,L,1), log_probs (B,L,1), v_loc (B,L,1)).
        ``u`` / ``z`` optionally inject the uniforms / normals of the stochastic branch."""
        if available_actions is None:
            # the reference crashes here (transformer_act.py:13-22, SURVEY 8b "Errors"); be explicit
            raise TypeError("available_actions is required by the discrete act path")
        if u is None and z is None and self.device.type == "cuda":
            oh = self._host_f32(obs, (-1, self.n_agent, self.obs_dim))
            ah = None if oh is None else self._host_f32(available_actions, (oh.shape[0], self.n_agent, self.action_dim))
            if oh is not None and ah is not None and oh.shape[0] >= self.HOST_SLICE_MIN:
                actions, logp, values = self._get_actions_host(oh, ah, deterministic, stride)
                if self.action_type == "Discrete":
                    actions = actions.long()
                actions, logp = self._out(actions, logp)
                return actions, logp, values
        obs = self._dev(obs, (-1, self.n_agent, self.obs_dim))
        B, L = obs.shape[0], self.n_agent
        ava = self._dev(available_actions, (B, L, self.action_dim))
        u = None if u is None else self._dev(u, (B, L))
        z = None if z is None else self._dev(z, (B, 2))
        values = torch.empty(B, L, 1, **self.tpdv)
        actions = torch.empty(B, L, 1, **self.tpdv)
        logp = torch.empty(B, L, 1, **self.tpdv)
        step = self.rollout_step
        self.rollout_step = (self.rollout_step + 1) & 0xFFFFFFFF
        with self._ctx():
            h = self.sync_weights()
            _cabi.check(_cabi.lib().matas_policy_get_actions(
                h, _cabi.ptr(obs), _cabi.ptr(ava), B, int(bool(deterministic)), int(stride), self.rollout_seed, step,
                self.env_offset, _cabi.ptr(u), _cabi.ptr(z), _cabi.ptr(values), _cabi.ptr(actions), _cabi.ptr(logp),
                _cabi.stream_ptr()))
        if self.action_type == "Discrete":
            actions = actions.long()          # transformer_act.py:135
        actions, logp = self._out(actions, logp)
        return actions, logp, values

    def forward(self, state, obs, action, available_actions=None, THOP_FLAG=False):
        """ma_transformer.py:257-295 -> (action_log (B,L,1), v_loc (B,L,1), entropy (B,L,1)).
        Differentiable w.r.t. the parameters once the backward kernels are built in."""
        from .autograd import evaluate
        obs = self._dev(obs, (-1, self.n_agent, self.obs_dim))
        B, L = obs.shape[0], self.n_agent
        action = self._dev(action, (B, L, 1))
        ava = None if available_actions is None else self._dev(available_actions, (B, L, self.action_dim))
        return evaluate(self, obs, action, ava)
