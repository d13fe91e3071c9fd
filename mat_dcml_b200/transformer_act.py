"""Drop-ins for the reference's second-level entry points
(mat_src/mat/algorithms/utils/transformer_act.py:30,103,131,176 -- spelling ``autoregreesive`` is
the reference's and part of the API).  Same positional signatures; ``decoder`` must be the
``Decoder`` of a mat_dcml_b200 ``MultiAgentTransformer`` (it carries the kernel handle), and
``obs_rep`` its encoder output.  ``obs``, ``batch_size``, ``action_dim`` and ``tpdv`` are accepted
for compatibility and, like in the reference, do not influence the result.
"""
import torch

from . import _cabi


def _model(decoder):
    owner = getattr(decoder, "_owner", None)
    if not owner:
        raise TypeError("decoder is not a mat_dcml_b200 Decoder (no kernel handle attached)")
    return owner[0]


def _autoregressive(decoder, obs_rep, n_agent, available_actions, deterministic, stride, mode, u=None, z=None):
    m = _model(decoder)
    want = _cabi.MODE_SEMI if m.action_type == "Semi_Discrete" else _cabi.MODE_DISCRETE
    if mode != want:
        raise TypeError("model was built for action_type %s" % m.action_type)
    if available_actions is None:
        raise TypeError("available_actions is required by the discrete act path")
    rep = m._dev(obs_rep, (-1, n_agent, m.n_embd))
    B = rep.shape[0]
    ava = m._dev(available_actions, (B, n_agent, m.action_dim))
    u = None if u is None else m._dev(u, (B, n_agent))
    z = None if z is None else m._dev(z, (B, 2))
    actions = torch.empty(B, n_agent, 1, **m.tpdv)
    logp = torch.empty(B, n_agent, 1, **m.tpdv)
    step = m.rollout_step
    m.rollout_step = (m.rollout_step + 1) & 0xFFFFFFFF
    with m._ctx():
        h = m.sync_weights()
        _cabi.check(_cabi.lib().matas_policy_act_autoregressive(
            h, _cabi.ptr(rep), _cabi.ptr(ava), B, int(bool(deterministic)), int(stride), m.rollout_seed, step,
            m.env_offset, _cabi.ptr(u), _cabi.ptr(z), _cabi.ptr(actions), _cabi.ptr(logp), _cabi.stream_ptr()))
    return actions, logp


def _parallel(decoder, obs_rep, action, n_agent, available_actions):
    m = _model(decoder)
    rep = m._dev(obs_rep, (-1, n_agent, m.n_embd))
    B = rep.shape[0]
    act = m._dev(action, (B, n_agent, 1))
    ava = None if available_actions is None else m._dev(available_actions, (B, n_agent, m.action_dim))
    logp = torch.empty(B, n_agent, 1, **m.tpdv)
    ent = torch.empty(B, n_agent, 1, **m.tpdv)
    with m._ctx():
        h = m.sync_weights()
        _cabi.check(_cabi.lib().matas_policy_act_parallel(h, _cabi.ptr(rep), _cabi.ptr(ava), _cabi.ptr(act), B,
                                                          _cabi.ptr(logp), _cabi.ptr(ent), _cabi.stream_ptr()))
    return logp, ent


def semi_discrete_autoregreesive_act(decoder, obs_rep, obs, batch_size, n_agent, action_dim, tpdv,
                                     available_actions=None, deterministic=False, semi_index=-1, stride=2):
    if semi_index != -1:
        raise NotImplementedError("semi_index must be -1")
    m = _model(decoder)
    a, lp = _autoregressive(decoder, obs_rep, n_agent, available_actions, deterministic, stride, _cabi.MODE_SEMI)
    return m._out(a, lp)


def semi_discrete_parallel_act(decoder, obs_rep, obs, action, batch_size, n_agent, action_dim, tpdv,
                               available_actions=None, semi_index=-1):
    if semi_index != -1:
        raise NotImplementedError("semi_index must be -1")
    return _parallel(decoder, obs_rep, action, n_agent, available_actions)


def discrete_autoregreesive_act(decoder, obs_rep, obs, batch_size, n_agent, action_dim, tpdv,
                                available_actions=None, deterministic=False):
    m = _model(decoder)
    a, lp = _autoregressive(decoder, obs_rep, n_agent, available_actions, deterministic, 4, _cabi.MODE_DISCRETE)
    return m._out(a.long(), lp)


def discrete_parallel_act(decoder, obs_rep, obs, action, batch_size, n_agent, action_dim, tpdv,
                          available_actions=None):
    return _parallel(decoder, obs_rep, action, n_agent, available_actions)
