"""CPU restatement of the MAT-AS policy forward (test infrastructure; see oracle/__init__.py).

Functional torch-CPU fp32 code over a plain ``state_dict`` with the reference's
parameter names.  It follows the reference's algorithm, including the things a
fast implementation would not do (no KV cache: the whole decoder is re-run for
every agent, ``transformer_act.py:77-98``), so that it doubles as the timed CPU
baseline.  Citations are to ``/root/reference/mat_src/mat/algorithms``.

Differences from the reference are confined to randomness: the stochastic
decode takes the uniforms / normals as inputs
instead of calling ``Categorical.sample`` / ``Normal.sample``.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

NORMAL_STD = 0.5          # utils/transformer_act.py:6
MASK_LOGIT = -1e10        # utils/transformer_act.py:14
HALF_LOG_2PI = 0.5 * math.log(2.0 * math.pi)


# --------------------------------------------------------------------------
# parameter inventory (mat/algorithm/ma_transformer.py:119-203)
# --------------------------------------------------------------------------
def param_table(n_agent, obs_dim, state_dim, act_dim=2, n_block=2, n_embd=64, semi=True):
    """Ordered (name, shape, kind) list of the trainable parameters, in ``state_dict`` order.
    kind: 'ln_w' | 'ln_b' | 'w_act' (Linear feeding a GELU) | 'w' | 'b' | 'log_std'.
    The ``attn*.mask`` buffers are not parameters; see ``mask_names``."""
    D = n_embd
    out = []

    def lin(prefix, o, i, act=False, bias=True):
        out.append((prefix + ".weight", (o, i), "w_act" if act else "w"))
        if bias:
            out.append((prefix + ".bias", (o,), "b"))

    def ln(prefix, d):
        out.append((prefix + ".weight", (d,), "ln_w"))
        out.append((prefix + ".bias", (d,), "ln_b"))

    def attn(prefix):
        for nm in ("key", "query", "value", "proj"):
            lin(prefix + "." + nm, D, D)

    ln("encoder.state_encoder.0", state_dim)
    lin("encoder.state_encoder.1", D, state_dim, act=True)
    ln("encoder.obs_encoder.0", obs_dim)
    lin("encoder.obs_encoder.1", D, obs_dim, act=True)
    ln("encoder.ln", D)
    for b in range(n_block):
        p = "encoder.blocks.%d" % b
        ln(p + ".ln1", D)
        ln(p + ".ln2", D)
        attn(p + ".attn")
        lin(p + ".mlp.0", D, D, act=True)
        lin(p + ".mlp.2", D, D)
    lin("encoder.head.0", D, D, act=True)
    ln("encoder.head.2", D)
    lin("encoder.head.3", 1, D)
    if semi:
        out.append(("decoder.log_std", (act_dim,), "log_std"))
    lin("decoder.action_encoder.0", D, act_dim + 1, act=True, bias=False)
    ln("decoder.obs_encoder.0", obs_dim)
    lin("decoder.obs_encoder.1", D, obs_dim, act=True)
    ln("decoder.ln", D)
    for b in range(n_block):
        p = "decoder.blocks.%d" % b
        ln(p + ".ln1", D)
        ln(p + ".ln2", D)
        ln(p + ".ln3", D)
        attn(p + ".attn1")
        attn(p + ".attn2")
        lin(p + ".mlp.0", D, D, act=True)
        lin(p + ".mlp.2", D, D)
    lin("decoder.head.0", D, D, act=True)
    ln("decoder.head.2", D)
    lin("decoder.head.3", act_dim, D)
    return out


def param_shapes(*a, **k):
    return [(n, s) for n, s, _ in param_table(*a, **k)]


def make_weights(seed, n_agent, obs_dim=7, state_dim=None, act_dim=2, profile="trained", semi=True):
    """Deterministic synthetic weights (numpy RandomState, stable across versions).

    ``profile='init'`` mimics the magnitudes of the reference initialisation
    (orthogonal, gain 0.01 for plain layers and sqrt(2) for activated ones,
    zero bias; ``ma_transformer.py:18-21``): logits ~1e-2.  ``'trained'`` adds
    O(0.3) noise everywhere, as SURVEY.md section 4 recommends, so that the
    top-2 logit gaps are well clear of the 1e-4 tie tolerance.
    """
    if state_dim is None:
        state_dim = n_agent + 1
    rng = np.random.RandomState(seed)
    noisy = profile == "trained"
    sd = {}
    for name, shape, kind in param_table(n_agent, obs_dim, state_dim, act_dim, semi=semi):
        x = rng.standard_normal(shape)
        if kind == "log_std":
            v = 1.0 + (0.3 * x if noisy else 0.0 * x)
        elif kind == "ln_w":
            v = 1.0 + (0.2 * x if noisy else 0.0 * x)
        elif kind in ("ln_b", "b"):
            v = (0.2 if kind == "ln_b" else 0.3) * x if noisy else 0.0 * x
        else:
            gain = math.sqrt(2.0) if kind == "w_act" else 0.01
            v = gain * x / math.sqrt(max(shape))
            if noisy:
                v = v + 0.6 * rng.standard_normal(shape) / math.sqrt(shape[1])
        sd[name] = torch.from_numpy(np.ascontiguousarray(v, dtype=np.float32))
    return sd


def mask_names(n_block=2):
    names = ["encoder.blocks.%d.attn.mask" % b for b in range(n_block)]
    for b in range(n_block):
        names += ["decoder.blocks.%d.attn1.mask" % b, "decoder.blocks.%d.attn2.mask" % b]
    return names


def full_state_dict(sd, n_agent, n_block=2):
    """Add the reference's tril ``mask`` buffers so the dict loads with strict=True."""
    out = dict(sd)
    m = torch.tril(torch.ones(n_agent + 1, n_agent + 1)).view(1, 1, n_agent + 1, n_agent + 1)
    for nm in mask_names(n_block):
        out[nm] = m.clone()
    return out


# --------------------------------------------------------------------------
# layers
# --------------------------------------------------------------------------
def _lin(sd, p, x):
    return F.linear(x, sd[p + ".weight"], sd.get(p + ".bias"))


def _ln(sd, p, x):
    return F.layer_norm(x, (x.shape[-1],), sd[p + ".weight"], sd[p + ".bias"], 1e-5)


def _attention(sd, p, key, value, query, causal, n_head=2):
    """SelfAttention.forward, ma_transformer.py:45-69."""
    B, L, D = query.shape
    hs = D // n_head

    def split(t):
        return t.view(B, L, n_head, hs).transpose(1, 2)

    k = split(_lin(sd, p + ".key", key))
    q = split(_lin(sd, p + ".query", query))
    v = split(_lin(sd, p + ".value", value))
    att = (q @ k.transpose(-2, -1)) * (1.0 / math.sqrt(hs))
    if causal:
        keep = torch.tril(torch.ones(L, L, dtype=torch.bool))
        att = att.masked_fill(~keep, float("-inf"))
    att = F.softmax(att, dim=-1)
    y = (att @ v).transpose(1, 2).contiguous().view(B, L, D)
    return _lin(sd, p + ".proj", y)


def encoder(sd, obs, n_block=2):
    """Encoder.forward with encode_state=False, ma_transformer.py:141-154 -> (v_loc, rep)."""
    x = F.gelu(_lin(sd, "encoder.obs_encoder.1", _ln(sd, "encoder.obs_encoder.0", obs)))
    x = _ln(sd, "encoder.ln", x)
    for b in range(n_block):
        p = "encoder.blocks.%d" % b
        x = _ln(sd, p + ".ln1", x + _attention(sd, p + ".attn", x, x, x, causal=False))
        x = _ln(sd, p + ".ln2", x + _lin(sd, p + ".mlp.2", F.gelu(_lin(sd, p + ".mlp.0", x))))
    h = _ln(sd, "encoder.head.2", F.gelu(_lin(sd, "encoder.head.0", x)))
    return _lin(sd, "encoder.head.3", h), x


def decoder(sd, shifted, rep, n_block=2):
    """Decoder.forward (dec_actor=False), ma_transformer.py:211-230, DecodeBlock :112-116."""
    x = _ln(sd, "decoder.ln", F.gelu(_lin(sd, "decoder.action_encoder.0", shifted)))
    for b in range(n_block):
        p = "decoder.blocks.%d" % b
        x = _ln(sd, p + ".ln1", x + _attention(sd, p + ".attn1", x, x, x, causal=True))
        x = _ln(sd, p + ".ln2", rep + _attention(sd, p + ".attn2", x, x, rep, causal=True))
        x = _ln(sd, p + ".ln3", x + _lin(sd, p + ".mlp.2", F.gelu(_lin(sd, p + ".mlp.0", x))))
    h = _ln(sd, "decoder.head.2", F.gelu(_lin(sd, "decoder.head.0", x)))
    return _lin(sd, "decoder.head.3", h)


# --------------------------------------------------------------------------
# action heads (utils/transformer_act.py:8-28)
# --------------------------------------------------------------------------
def _masked_logz(logit, ava):
    """Mask with -1e10 then normalise by logsumexp (what Categorical(logits=) does)."""
    if ava is not None:
        logit = torch.where(ava == 0, torch.full_like(logit, MASK_LOGIT), logit)
    return logit - torch.logsumexp(logit, dim=-1, keepdim=True)


def _pick_discrete(logit, ava, deterministic, u):
    z = _masked_logz(logit, ava)
    p = torch.softmax(z, dim=-1)
    if deterministic:
        a = p.argmax(dim=-1)
    else:
        # inverse-CDF on the injected uniform; for two classes: a = [u >= p0]
        cdf = torch.cumsum(p, dim=-1)
        a = (u.unsqueeze(-1) >= cdf[..., :-1]).sum(dim=-1)
    return a, z.gather(-1, a.unsqueeze(-1)).squeeze(-1)


def action_std(sd):
    return torch.sigmoid(sd["decoder.log_std"]) * NORMAL_STD


def _normal_logp(a, mean, std):
    return -((a - mean) ** 2) / (2.0 * std * std) - torch.log(std) - HALF_LOG_2PI


def _pick_continuous(mean, std, deterministic, z):
    a = mean if deterministic else mean + std * z
    return a, _normal_logp(a, mean, std)


# --------------------------------------------------------------------------
# autoregressive act
# --------------------------------------------------------------------------
def semi_discrete_autoregressive_act(sd, rep, ava, deterministic=False, stride=2, u=None, z=None,
                                     semi_index=-1):
    """utils/transformer_act.py:30-100.  ``u`` (B,L) uniforms and ``z`` (B,2) normals are the
    injected randomness of the stochastic branch."""
    B, L, _ = rep.shape
    A = 2
    nd = L + semi_index
    shifted = torch.zeros(B, L, A + 1)
    shifted[:, 0, 0] = 1.0
    act = torch.zeros(B, L, 1)
    logp = torch.zeros(B, L, 1)
    std = action_std(sd)

    def emit(j, lg):
        if j < nd:
            a, lp = _pick_discrete(lg, ava[:, j, :], deterministic, None if u is None else u[:, j])
            act[:, j, 0] = a.to(torch.float32)
            logp[:, j, 0] = lp
            if j + 1 < L:
                shifted[:, j + 1, 1:] = F.one_hot(a, A).to(torch.float32)
        else:
            a, lp = _pick_continuous(lg, std, deterministic, z)
            act[:, j, 0] = a[:, 1]
            logp[:, j, 0] = lp[:, 1]
            if j + 1 < L:
                shifted[:, j + 1, :] = a

    if not deterministic:
        for j in range(L):
            emit(j, decoder(sd, shifted, rep)[:, j, :])
        return act, logp

    s, e, go = 0, 1, True
    while e <= L and go:
        if e == L:
            go = False
        lg = decoder(sd, shifted, rep)[:, s:e, :]
        for j in range(s, e):
            emit(j, lg[:, j - s, :])
        if e < nd:
            s, e = e, min(e + stride, nd)
        else:
            s, e = e, min(e + 1, L)
    return act, logp


def discrete_autoregressive_act(sd, rep, ava, deterministic=False, u=None):
    """utils/transformer_act.py:131-173 (incl. the stride-4 / own-row one-hot quirk of the
    deterministic branch, :153,156)."""
    B, L, _ = rep.shape
    A = 2
    shifted = torch.zeros(B, L, A + 1)
    shifted[:, 0, 0] = 1.0
    act = torch.zeros(B, L, 1, dtype=torch.long)
    logp = torch.zeros(B, L, 1)
    if deterministic:
        s, e, go = 0, 1, True
        while e <= L and go:
            if e == L:
                go = False
            lg = decoder(sd, shifted, rep)[:, s:e, :]
            for j in range(s, e):
                a, lp = _pick_discrete(lg[:, j - s, :], ava[:, j, :], True, None)
                act[:, j, 0] = a
                logp[:, j, 0] = lp
                if (j - s) + 1 < L:
                    shifted[:, j, 1:] = F.one_hot(a, A).to(torch.float32)
            if e < L:
                s, e = e, min(e + 4, L)
    else:
        for j in range(L):
            lg = decoder(sd, shifted, rep)[:, j, :]
            a, lp = _pick_discrete(lg, None if ava is None else ava[:, j, :], False, u[:, j])
            act[:, j, 0] = a
            logp[:, j, 0] = lp
            if j + 1 < L:
                shifted[:, j + 1, 1:] = F.one_hot(a, A).to(torch.float32)
    return act, logp


# --------------------------------------------------------------------------
# teacher-forced (parallel) act
# --------------------------------------------------------------------------
def _cat_entropy(z):
    zc = torch.clamp(z, min=torch.finfo(z.dtype).min)
    return -(torch.exp(zc) * zc).sum(-1)


def semi_discrete_parallel_act(sd, rep, action, ava, semi_index=-1):
    """utils/transformer_act.py:103-129 -> (log_prob (B,L,1), entropy (B,L,1))."""
    B, L, _ = rep.shape
    A = 2
    nd = L + semi_index
    disc = action[:, :nd, 0].to(torch.long)
    shifted = torch.zeros(B, L, A + 1)
    shifted[:, 0, 0] = 1.0
    both = torch.cat([F.one_hot(disc, A).to(torch.float32), action[:, nd:, :].expand(-1, -1, A)], dim=1)
    shifted[:, 1:, 1:] = both[:, :-1, :]
    lg = decoder(sd, shifted, rep)
    z = _masked_logz(lg[:, :nd, :], None if ava is None else ava[:, :nd, :])
    lp_d = z.gather(-1, disc.unsqueeze(-1))
    ent_d = _cat_entropy(z).unsqueeze(-1)
    std = action_std(sd)
    lp_c = _normal_logp(action[:, nd:, :], lg[:, nd:, :], std)[:, :, -1:]
    ent_c = (0.5 + HALF_LOG_2PI + torch.log(std)).expand(B, L - nd, A)[:, :, -1:]
    return torch.cat([lp_d, lp_c], 1), torch.cat([ent_d, ent_c], 1)


def discrete_parallel_act(sd, rep, action, ava):
    """utils/transformer_act.py:176-189."""
    B, L, _ = rep.shape
    A = 2
    a = action[:, :, 0].to(torch.long)
    shifted = torch.zeros(B, L, A + 1)
    shifted[:, 0, 0] = 1.0
    shifted[:, 1:, 1:] = F.one_hot(a, A).to(torch.float32)[:, :-1, :]
    z = _masked_logz(decoder(sd, shifted, rep), ava)
    return z.gather(-1, a.unsqueeze(-1)), _cat_entropy(z).unsqueeze(-1)


# --------------------------------------------------------------------------
# policy-level entry points (algorithm/transformer_policy.py:116-217)
# --------------------------------------------------------------------------
def _t(x):
    return torch.as_tensor(np.asarray(x) if not torch.is_tensor(x) else x).to(torch.float32)


@torch.no_grad()
def get_actions(sd, obs, ava, n_agent, deterministic=False, stride=1, u=None, z=None, mode="semi"):
    obs = _t(obs).reshape(-1, n_agent, obs.shape[-1])
    ava = None if ava is None else _t(ava).reshape(-1, n_agent, 2)
    v, rep = encoder(sd, obs)
    if mode == "semi":
        a, lp = semi_discrete_autoregressive_act(sd, rep, ava, deterministic, stride,
                                                 None if u is None else _t(u), None if z is None else _t(z))
    else:
        a, lp = discrete_autoregressive_act(sd, rep, ava, deterministic, None if u is None else _t(u))
    return v.reshape(-1, 1), a.reshape(-1, 1), lp.reshape(-1, 1)


@torch.no_grad()
def get_values(sd, obs, n_agent):
    obs = _t(obs).reshape(-1, n_agent, obs.shape[-1])
    return encoder(sd, obs)[0].reshape(-1, 1)


def evaluate_actions(sd, obs, actions, ava, n_agent, active_masks=None, mode="semi"):
    """-> values (B*L,1), log_probs (B*L,1), entropy scalar (transformer_policy.py:181-217).
    Differentiable w.r.t. the tensors in ``sd`` (used for gradient parity)."""
    obs = _t(obs).reshape(-1, n_agent, obs.shape[-1])
    actions = _t(actions).reshape(-1, n_agent, 1)
    ava = None if ava is None else _t(ava).reshape(-1, n_agent, 2)
    v, rep = encoder(sd, obs)
    if mode == "semi":
        lp, ent = semi_discrete_parallel_act(sd, rep, actions, ava)
    else:
        lp, ent = discrete_parallel_act(sd, rep, actions, ava)
    lp, ent = lp.reshape(-1, 1), ent.reshape(-1, 1)
    if active_masks is not None:
        am = _t(active_masks).reshape(-1, 1)
        ent_s = (ent * am).sum() / am.sum()
    else:
        ent_s = ent.mean()
    return v.reshape(-1, 1), lp, ent_s
