"""Diagonal Gaussian posterior of the VAE encoder (stable_diffusion/modules/distributions.py:5-19).
Out of the hot path: plain tensor arithmetic on whatever device the moments live on."""
import torch


class GaussianDistribution:
    def __init__(self, moments: torch.Tensor):
        self.mean, self.log_var = torch.chunk(moments, 2, dim=1)

    def sample(self, eps=None):
        """``eps`` (extension): the N(0, 1) draw to use instead of ``torch.randn_like(std)`` (seeded parity tests)."""
        std = torch.exp(0.5 * self.log_var)
        return self.mean + (torch.randn_like(std) if eps is None else eps) * std

    def kl(self):
        self.var = torch.exp(self.log_var)
        return 0.5 * torch.sum(torch.pow(self.mean, 2) + self.var - 1.0 - self.log_var, dim=[1, 2, 3])
