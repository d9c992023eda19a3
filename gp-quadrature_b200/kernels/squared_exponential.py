"""Squared-exponential kernel k(r) = variance * exp(-r^2 / (2 l^2)) and its spectral density
(mirror of the reference's kernels/squared_exponential.py:9-216)."""
import math

import torch

from .kernel import Kernel


class SquaredExponential(Kernel):
    _median_scale = 0.5  # the reference halves the median distance for SE (squared_exponential.py:207)

    def kernel(self, distance):
        ell, var = self.get_hyper("lengthscale"), self.get_hyper("variance")
        return var * torch.exp(-0.5 * (distance / ell) ** 2)

    def _kernel_scalar(self, r):
        ell, var = self.get_hyper("lengthscale"), self.get_hyper("variance")
        return var * math.exp(-0.5 * (r / ell) ** 2)

    def _scalar_closures(self):
        """(k(r), S(r)) on python floats with the hypers read once (used by utils.kernels.get_xis)."""
        ell, var, d = self.get_hyper("lengthscale"), self.get_hyper("variance"), self.dimension
        pref = ((2 * math.pi) * ell ** 2) ** (d / 2) * var
        c = ((2 * math.pi) ** 2) * ell ** 2
        return (lambda r: var * math.exp(-0.5 * (r / ell) ** 2)), (lambda r: pref * math.exp(-c * (r * r) / 2))

    def _spectral_family(self):
        """(family, c, a, b, p) with S(xi) = c exp(-a |xi|^2): what efgp_fit evaluates on the device."""
        ell, var, d = self.get_hyper("lengthscale"), self.get_hyper("variance"), self.dimension
        return 0, ((2 * math.pi) * ell ** 2) ** (d / 2) * var, ((2 * math.pi) ** 2) * ell ** 2 / 2, 0.0, 0.0

    def _sq_norm(self, xid):
        if xid.ndim == 1:
            xid = xid.unsqueeze(-1)
        return torch.sum(xid ** 2, dim=-1)

    def spectral_density(self, xid):
        """S(xi) = (2 pi l^2)^(d/2) variance exp(-2 pi^2 l^2 |xi|^2)  (squared_exponential.py:64-91)."""
        ell, var = self.get_hyper("lengthscale"), self.get_hyper("variance")
        q = self._sq_norm(xid)
        pref = ((2 * math.pi) * ell ** 2) ** (self.dimension / 2) * var
        return pref * torch.exp(-((2 * math.pi) ** 2) * ell ** 2 * q / 2)

    def _spectral_density_scalar(self, r):
        ell, var = self.get_hyper("lengthscale"), self.get_hyper("variance")
        pref = ((2 * math.pi) * ell ** 2) ** (self.dimension / 2) * var
        return pref * math.exp(-((2 * math.pi) ** 2) * ell ** 2 * (r * r) / 2)

    def spectral_grad(self, xid):
        """[dS/dl, dS/dvariance]  (squared_exponential.py:93-123)."""
        ell, var = self.get_hyper("lengthscale"), self.get_hyper("variance")
        q = self._sq_norm(xid)
        two_pi_sq = (2 * math.pi) ** 2
        pref = ((2 * math.pi) * ell ** 2) ** (self.dimension / 2) * var
        s = pref * torch.exp(-two_pi_sq * ell ** 2 * q / 2)
        return torch.stack([s * (self.dimension / ell - two_pi_sq * ell * q), s / var], dim=-1)
