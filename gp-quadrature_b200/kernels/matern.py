"""Matern kernel (nu = 1/2, 3/2, 5/2 in closed form) and its spectral density
(mirror of the reference's kernels/matern.py:9-264)."""
import math

import torch

from .kernel import Kernel


class Matern(Kernel):
    _median_scale = 1.0

    def __init__(self, dimension, init_lengthscale=float("nan"), init_variance=float("nan"), nu=2.5, **extra):
        if not (0.1 <= float(nu) <= 10.0):
            raise ValueError("nu must be in [0.1, 10]")
        self.nu = float(nu)
        super().__init__(dimension, init_lengthscale, init_variance, **extra)

    def kernel(self, distance):
        ell, var = self.lengthscale, self.variance
        s = torch.abs(distance) / ell
        if self.nu == 0.5:
            return var * torch.exp(-s)
        if self.nu == 1.5:
            return var * (1 + math.sqrt(3) * s) * torch.exp(-math.sqrt(3) * s)
        if self.nu == 2.5:
            return var * (1 + math.sqrt(5) * s + 5 * s ** 2 / 3) * torch.exp(-math.sqrt(5) * s)
        # general nu through the modified Bessel function (the reference's branch, matern.py:72-98,
        # references undefined names and cannot run; this is the formula its docstring states)
        from scipy import special
        z = torch.clamp(math.sqrt(2.0 * self.nu) * torch.abs(distance) / ell, min=1e-8)
        kv = torch.as_tensor(special.kv(self.nu, z.cpu().numpy()), device=distance.device, dtype=distance.dtype)
        K = var * 2 ** (1 - self.nu) / special.gamma(self.nu) * z ** self.nu * kv
        return torch.nan_to_num(K, nan=var, posinf=var, neginf=0.0)

    def _kernel_scalar(self, r):
        ell, var = self.lengthscale, self.variance
        s = abs(r) / ell
        if self.nu == 0.5:
            return var * math.exp(-s)
        if self.nu == 1.5:
            return var * (1 + math.sqrt(3) * s) * math.exp(-math.sqrt(3) * s)
        if self.nu == 2.5:
            return var * (1 + math.sqrt(5) * s + 5 * s ** 2 / 3) * math.exp(-math.sqrt(5) * s)
        return float(self.kernel(torch.tensor(r, dtype=torch.float64)))

    def _scalar_closures(self):
        ell, var, nu, d = self.lengthscale, self.variance, self.nu, self.dimension
        sc = self._scaling(ell)
        if nu == 0.5:
            kf = lambda r: var * math.exp(-(abs(r) / ell))
        elif nu == 1.5:
            kf = lambda r: var * (1 + math.sqrt(3) * (abs(r) / ell)) * math.exp(-math.sqrt(3) * (abs(r) / ell))
        elif nu == 2.5:
            kf = lambda r: (var * (1 + math.sqrt(5) * (abs(r) / ell) + 5 * (abs(r) / ell) ** 2 / 3)
                            * math.exp(-math.sqrt(5) * (abs(r) / ell)))
        else:
            kf = self._kernel_scalar
        sf = lambda r: var * sc * (2 * nu / ell ** 2 + (4 * math.pi ** 2) * (r * r)) ** (-(nu + d / 2))
        return kf, sf

    def _spectral_family(self):
        """(family, c, a, b, p) with S(xi) = c (b + 4 pi^2 |xi|^2)^(-p): what efgp_fit evaluates on the device."""
        ell, var, nu, d = self.lengthscale, self.variance, self.nu, self.dimension
        return 1, var * self._scaling(ell), 0.0, 2 * nu / ell ** 2, nu + d / 2

    def _scaling(self, ell):
        d, nu = self.dimension, self.nu
        return ((2 * math.sqrt(math.pi)) ** d * math.gamma(nu + d / 2) * (2 * nu) ** nu
                / (math.gamma(nu) * ell ** (2 * nu)))

    def spectral_density(self, xid):
        """S = variance c(l) (2 nu / l^2 + 4 pi^2 |xi|^2)^-(nu + d/2)  (matern.py:100-122)."""
        if xid.ndim == 1:
            xid = xid.unsqueeze(-1)
        q = torch.sum(xid ** 2, dim=-1)
        ell, var = self.lengthscale, self.variance
        return var * self._scaling(ell) * (2 * self.nu / ell ** 2 + (4 * math.pi ** 2) * q) ** (
            -(self.nu + self.dimension / 2))

    def _spectral_density_scalar(self, r):
        ell, var = self.lengthscale, self.variance
        return var * self._scaling(ell) * (2 * self.nu / ell ** 2 + (4 * math.pi ** 2) * (r * r)) ** (
            -(self.nu + self.dimension / 2))

    def spectral_grad(self, xid):
        """[dS/dl, dS/dvariance]  (matern.py:124-168)."""
        if xid.ndim == 1:
            xid = xid.unsqueeze(-1)
        S = self.spectral_density(xid)
        q = torch.sum(xid ** 2, dim=-1)
        ell, var, nu = self.lengthscale, self.variance, self.nu
        den = 2 * nu / ell ** 2 + (4 * math.pi ** 2) * q
        power = -(nu + self.dimension / 2)
        dl = S * (-2 * nu / ell + power * (-4 * nu / ell ** 3) / den)
        return torch.stack([dl, S / var], dim=-1)
