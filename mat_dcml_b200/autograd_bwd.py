"""autograd bridge for MultiAgentTransformer.forward: forward = matas_policy_evaluate_train,
backward = matas_policy_evaluate_backward, which accumulates straight into the model's flat
gradient buffer (every parameter's .grad is a view of it), so nothing is returned per parameter."""
import torch

from . import _cabi


class EvaluateFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, model, obs, action, ava, anchor):
        B, L = obs.shape[0], model.n_agent
        values = torch.empty(B, L, 1, **model.tpdv)
        logp = torch.empty(B, L, 1, **model.tpdv)
        ent = torch.empty(B, L, 1, **model.tpdv)
        with model._ctx():
            h = model.sync_weights()
            _cabi.check(_cabi.lib().matas_policy_evaluate_train(h, _cabi.ptr(obs), _cabi.ptr(ava), _cabi.ptr(action), B,
                                                                _cabi.ptr(values), _cabi.ptr(logp), _cabi.ptr(ent),
                                                                _cabi.stream_ptr()))
        ctx.model, ctx.B = model, B
        ctx.save_for_backward(obs, action, ava if ava is not None else torch.empty(0, device=obs.device))
        ctx.has_ava = ava is not None
        return logp, values, ent

    @staticmethod
    def backward(ctx, d_logp, d_values, d_ent):
        model = ctx.model
        obs, action, ava = ctx.saved_tensors
        z = lambda g, like: (torch.zeros(ctx.B, model.n_agent, 1, **model.tpdv) if g is None else g.contiguous().float())
        d_logp, d_values, d_ent = z(d_logp, None), z(d_values, None), z(d_ent, None)
        grad = model.flat_grad()
        with model._ctx():
            _cabi.check(_cabi.lib().matas_policy_evaluate_backward(
                model._h(), _cabi.ptr(obs), _cabi.ptr(ava) if ctx.has_ava else None, _cabi.ptr(action), ctx.B,
                _cabi.ptr(d_values), _cabi.ptr(d_logp), _cabi.ptr(d_ent), _cabi.ptr(grad), _cabi.stream_ptr()))
        return None, None, None, None, None
