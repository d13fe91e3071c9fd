"""ValueNorm (mat_src/mat/utils/valuenorm.py:8-80, input_shape=1) with its three running scalars
in one device tensor so that the GAE kernel can read them without a host sync."""
import numpy as np
import torch


class ValueNorm:
    def __init__(self, input_shape=1, norm_axes=1, beta=0.99999, per_element_update=False, epsilon=1e-5,
                 device=torch.device("cpu")):
        assert input_shape == 1 and norm_axes == 1, "MAT-AS uses a scalar value normaliser (mat_trainer.py:49-52)"
        self.beta, self.epsilon, self.per_element_update = beta, epsilon, per_element_update
        self.tpdv = dict(dtype=torch.float32, device=torch.device(device))
        self.stats = torch.zeros(3, **self.tpdv)      # running_mean, running_mean_sq, debiasing_term
        self.updated = False

    running_mean = property(lambda s: s.stats[0:1])
    running_mean_sq = property(lambda s: s.stats[1:2])
    debiasing_term = property(lambda s: s.stats[2])

    def running_mean_var(self):
        d = self.stats[2].clamp(min=self.epsilon)
        mean = self.stats[0:1] / d
        var = (self.stats[1:2] / d - mean ** 2).clamp(min=1e-2)
        return mean, var

    def _t(self, x):
        if isinstance(x, np.ndarray):
            x = torch.from_numpy(x)
        return x.to(**self.tpdv)

    @torch.no_grad()
    def update(self, input_vector):
        self.updated = True
        x = self._t(input_vector)
        bm, bsq = x.mean(dim=0), (x ** 2).mean(dim=0)
        w = self.beta ** int(np.prod(x.size()[:1])) if self.per_element_update else self.beta
        self.stats[0:1].mul_(w).add_(bm * (1.0 - w))
        self.stats[1:2].mul_(w).add_(bsq * (1.0 - w))
        self.stats[2].mul_(w).add_(1.0 * (1.0 - w))

    def normalize(self, input_vector):
        mean, var = self.running_mean_var()
        return (self._t(input_vector) - mean) / torch.sqrt(var)

    def denormalize(self, input_vector):
        mean, var = self.running_mean_var()
        return (self._t(input_vector) * torch.sqrt(var) + mean).cpu().numpy()
