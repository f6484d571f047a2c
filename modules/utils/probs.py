"""Gaussian prior log-probability -- the part of the reference's modules/utils/probs.py (40-102) the
discrete model touches: GaussianDistribution.standard_logp (:87-93) feeds logp_x in
PointInterpFlow.log_prob (interpflow.py:343-344).  Unlike the reference it does not allocate on a
hard-coded 'cuda:0' at construction (probs.py:51,56)."""
import numpy as np
import torch
from torch import Tensor


class GaussianDistribution(object):
    Log2PI = float(np.log(2 * np.pi))

    def __init__(self, pc_channel: int, mu: float, vars: float, temperature: float = 1.0, device=None):
        assert temperature >= 0.0 and temperature <= 1.0
        self.pc_channel, self.mu, self.vars = pc_channel, mu, vars
        self.temperature = temperature * temperature

    @staticmethod
    def standard_likelood(x):
        return -0.5 * (x ** 2 + GaussianDistribution.Log2PI)

    def standard_logp(self, z: Tensor):
        logp_z = GaussianDistribution.standard_likelood(z)
        sum_dims = tuple([i for i in range(1, logp_z.dim())])
        return torch.sum(logp_z, dim=sum_dims)   # [B,]

    def standard_sample(self, shape, device, temperature=None):
        temp = temperature ** 2 if temperature is not None else self.temperature
        return torch.normal(mean=torch.zeros(shape), std=(torch.ones(shape) * temp)).to(device)
