"""Kernel base class: hyperparameter plumbing shared by SquaredExponential and Matern.

Mirrors the observable behaviour of the reference's kernels/kernel.py:6-240 (constructor kwargs
`dimension`, `init_lengthscale`, `init_variance`; get_hyper / set_hyper / iter_hypers; the
GPParams back-reference created at construction) without pydantic.
"""
import math

import torch

from .kernel_params import GPParams


class Kernel:
    hypers = ["lengthscale", "variance"]
    num_hypers = 3  # kernel hypers + noise variance

    def __init__(self, dimension, init_lengthscale=float("nan"), init_variance=float("nan"), **extra):
        if int(dimension) < 1:
            raise ValueError("dimension must be >= 1")
        for name, v in (("init_lengthscale", init_lengthscale), ("init_variance", init_variance)):
            if not math.isnan(float(v)) and float(v) < 1e-6:
                raise ValueError(f"{name} must be >= 1e-6")
        self.dimension = int(dimension)
        self.init_lengthscale = float(init_lengthscale)
        self.init_variance = float(init_variance)
        self.hypers = list(type(self).hypers)
        self.num_hypers = type(self).num_hypers
        self._gp_params_ref = None
        self._params_dict = {"lengthscale": self.init_lengthscale, "variance": self.init_variance}
        # the reference always attaches a GPParams (noise 0.1) right after construction (kernel.py:66-95)
        GPParams(kernel=self, init_sig2=0.1)

    # ------------------------------------------------------------------ hyper access
    def get_hyper(self, name):
        """exp(raw[idx]) of the attached GPParams (reference kernel.py:97-119)."""
        ref = self._gp_params_ref
        if ref is None:
            if name in self._params_dict:
                return self._params_dict[name]
            raise RuntimeError(f"No GPParams reference available and unknown parameter: {name}")
        if name in ref.hypers_names:
            return float(ref.pos[ref.hypers_names.index(name)].item())
        raise ValueError(f"Unknown hyperparameter: {name}")

    def set_hyper(self, name, value):
        """Stores log(float32(value)) into raw (the float32 round trip of reference kernel.py:121-139)."""
        if name not in self.hypers:
            raise ValueError(f"Unknown hyperparameter: {name}")
        self._params_dict[name] = float(value)
        ref = self._gp_params_ref
        if ref is not None and name in ref.hypers_names:
            with torch.no_grad():
                ref.raw.data[ref.hypers_names.index(name)] = torch.log(torch.tensor(float(value)))

    def iter_hypers(self):
        for name in self.hypers:
            yield name, self._params_dict[name]

    @property
    def lengthscale(self):
        return self.get_hyper("lengthscale")

    @lengthscale.setter
    def lengthscale(self, v):
        self.set_hyper("lengthscale", v)

    @property
    def variance(self):
        return self.get_hyper("variance")

    @variance.setter
    def variance(self, v):
        self.set_hyper("variance", v)

    # ------------------------------------------------------------------ to be provided by subclasses
    def kernel(self, distance):
        raise NotImplementedError

    def spectral_density(self, xid):
        raise NotImplementedError

    def spectral_grad(self, xid):
        raise NotImplementedError

    # scalar (python float) versions used by the eps-driven grid search
    def _kernel_scalar(self, r):
        raise NotImplementedError

    def _spectral_density_scalar(self, r):
        raise NotImplementedError

    # ------------------------------------------------------------------ dense helpers (small N only)
    def kernel_matrix(self, x, y):
        if x.ndim == 1:
            x = x.unsqueeze(-1)
        if y.ndim == 1:
            y = y.unsqueeze(-1)
        return self.kernel(torch.cdist(x, y))

    def log_marginal(self, x, y, sigmasq):
        """Dense Cholesky log marginal likelihood (reference squared_exponential.py:125-160)."""
        if x.ndim == 1:
            x = x.unsqueeze(-1)
        n = x.shape[0]
        K = self.kernel_matrix(x, x) + sigmasq * torch.eye(n, device=x.device, dtype=x.dtype)
        try:
            L = torch.linalg.cholesky(K)
        except RuntimeError:
            return float("-inf")
        a = torch.cholesky_solve(y.unsqueeze(-1), L).squeeze(-1)
        return -(0.5 * torch.sum(y * a) + torch.sum(torch.log(torch.diag(L)))
                 + 0.5 * n * math.log(2 * math.pi)).item()

    _median_scale = 1.0

    def estimate_hyperparameters(self, x, y, K=1000):
        """Median-distance heuristic (reference squared_exponential.py:162-216, matern.py:212-264).
        Consumes the global torch RNG (randperm) when n > K, like the reference."""
        if x.ndim == 1:
            x = x.unsqueeze(-1)
        n = x.shape[0]
        y_var = torch.var(y).item()
        xs = x[torch.randperm(n)[:K].to(x.device)] if n > K else x
        d = torch.cdist(xs, xs)
        mask = d > 0
        med = torch.median(d[mask]).item() if bool(mask.any()) else 1.0
        return self._median_scale * med, y_var, 0.2 * y_var
