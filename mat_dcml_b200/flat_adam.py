"""``FlatAdam``: the policy's optimizer when the model lives on a GPU.

The reference builds ``torch.optim.Adam(self.transformer.parameters(), lr, eps, weight_decay)``
(mat_src/mat/algorithms/mat/algorithm/transformer_policy.py:103-105) and ``MATTrainer.ppo_update`` runs
``clip_grad_norm_`` + ``optimizer.step()`` over ~110 parameter tensors (mat_trainer.py:149-154).  Here every
parameter and gradient is a view into one flat fp32 buffer, so the same update is two kernel launches
(``matas_adam_clip_step``, csrc/adam.cu).  The class keeps what the reference's callers touch:
``param_groups[i]["lr"]`` (``lr_decay``, mat/utils/util.py:17-21), ``zero_grad()``, ``step()``, ``state_dict()``.
"""
import torch

from . import _cabi


class FlatAdam(torch.optim.Optimizer):
    def __init__(self, model, lr=5e-4, betas=(0.9, 0.999), eps=1e-5, weight_decay=0.0):
        super().__init__(list(model.parameters()), dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self._model = model
        flat = model.flat_parameters()
        if flat.device.type != "cuda":
            raise _cabi.MatasError("FlatAdam needs the model on a CUDA device (there is no CPU fallback)")
        self.exp_avg = torch.zeros_like(flat)
        self.exp_avg_sq = torch.zeros_like(flat)
        self.norm = torch.zeros(2, dtype=torch.float32, device=flat.device)     # {gradient norm, clip coefficient}
        self.step_count = 0
        # encoder.state_encoder.* and decoder.obs_encoder.* never receive a gradient (encode_state / dec_actor are off):
        # torch.optim.Adam skips parameters whose .grad is None, so weight decay must not touch them either
        self._frozen = [p for n, p in model.named_parameters()
                        if n.startswith("encoder.state_encoder.") or n.startswith("decoder.obs_encoder.")]

    def zero_grad(self, set_to_none=False):
        self._model.flat_grad(zero=True)

    @torch.no_grad()
    def clip_and_step(self, max_grad_norm=0.0):
        """clip_grad_norm_(parameters, max_grad_norm) + Adam step; returns the pre-clip gradient norm (device scalar)."""
        m = self._model
        g = self.param_groups[0]
        self.step_count += 1
        keep = [p.detach().clone() for p in self._frozen] if g["weight_decay"] != 0 else None
        with torch.cuda.device(m.device):
            _cabi.check(_cabi.lib().matas_adam_clip_step(
                _cabi.ptr(m.flat_parameters()), _cabi.ptr(m.flat_grad()), _cabi.ptr(self.exp_avg), _cabi.ptr(self.exp_avg_sq),
                m.flat_parameters().numel(), float(g["lr"]), float(g["betas"][0]), float(g["betas"][1]), float(g["eps"]),
                float(g["weight_decay"]), float(max_grad_norm or 0.0), self.step_count, _cabi.ptr(self.norm),
                _cabi.stream_ptr()))
        if keep is not None:
            for p, k in zip(self._frozen, keep):
                p.data.copy_(k)
        m.invalidate_weights()              # the kernel wrote the parameters behind torch's version counters
        return self.norm[0].clone()

    def step(self, closure=None):
        loss = closure() if closure is not None else None
        self.clip_and_step(0.0)
        return loss

    def state_dict(self):
        return {"step": self.step_count, "exp_avg": self.exp_avg.clone(), "exp_avg_sq": self.exp_avg_sq.clone(),
                "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups]}

    def load_state_dict(self, sd):
        self.step_count = int(sd["step"])
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        for g, s in zip(self.param_groups, sd["param_groups"]):
            g.update(s)
