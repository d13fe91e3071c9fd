"""Teacher-forced evaluation (MultiAgentTransformer.forward) over the C-ABI kernels.  Without grad
mode: matas_policy_evaluate.  With grad mode: a torch.autograd.Function whose backward is the
hand-written CUDA backward (autograd_bwd.py); there is no eager fallback."""
import torch

from . import _cabi


def _forward(model, obs, action, ava):
    B, L = obs.shape[0], model.n_agent
    values = torch.empty(B, L, 1, **model.tpdv)
    logp = torch.empty(B, L, 1, **model.tpdv)
    ent = torch.empty(B, L, 1, **model.tpdv)
    with model._ctx():
        h = model.sync_weights()
        _cabi.check(_cabi.lib().matas_policy_evaluate(h, _cabi.ptr(obs), _cabi.ptr(ava), _cabi.ptr(action), B,
                                                      _cabi.ptr(values), _cabi.ptr(logp), _cabi.ptr(ent),
                                                      _cabi.stream_ptr()))
    return logp, values, ent


def evaluate(model, obs, action, ava):
    """-> (action_log (B,L,1), v_loc (B,L,1), entropy (B,L,1))."""
    needs_grad = torch.is_grad_enabled() and any(p.requires_grad for p in model.parameters())
    if not needs_grad:
        return _forward(model, obs, action, ava)
    from .autograd_bwd import EvaluateFn
    return EvaluateFn.apply(model, obs, action, ava, model._anchor())
