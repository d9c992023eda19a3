"""Eps-driven equispaced frequency grid (mirror of the reference's utils/kernels.py:7-143).

`get_xis` returns (xis_1d, h, mtot): h = 1/(L + Ltime) with Ltime the eps-support of k(r), and
mtot = 2*ceil(Lfreq/h)+1 with Lfreq the eps-support of |r|^(d-1) S(r)/S(0).  The reference runs the
two 200-step bisections on 0-d torch tensors (~40 ms of host time per call, which would dominate a
GPU step); here the same bisection runs on Python floats through the kernels' scalar closed forms.
"""
import math

import torch


class GetTruncationBound:
    """Bisection for f(L) = eps with f monotonically decreasing (utils/kernels.py:7-69): the bracket
    [0, 1000] is doubled up to 10 times, then 200 halvings; the LAST MIDPOINT is returned."""

    def __init__(self, eps, kern, initial_upper_bound=1000.0, initial_lower_bound=0.0, max_iterations=200,
                 dtype=torch.float64):
        self.eps = eps
        self.kern = kern
        self.initial_upper_bound = initial_upper_bound
        self.initial_lower_bound = initial_lower_bound
        self.max_iterations = max_iterations
        self.dtype = dtype
        self._scalar = False

    def _f(self, v):
        if self._scalar:
            return self.kern(v)
        return self.kern(torch.tensor(v, device="cpu", dtype=self.dtype))

    def find_upper_bound_for_bisection(self):
        b = self.initial_upper_bound
        for _ in range(10):
            if self._f(b) > self.eps:
                b *= 2
            else:
                break
        return b

    def find_truncation_bound(self):
        a = self.initial_lower_bound
        b = self.find_upper_bound_for_bisection()
        mid = (a + b) / 2
        for _ in range(self.max_iterations):
            mid = (a + b) / 2
            if mid == a or mid == b:
                break          # interval collapsed to adjacent floats: every later midpoint is this one
            if self._f(mid) > self.eps:
                a = mid
            else:
                b = mid
        return mid


def _scalar_bound(eps, f):
    g = GetTruncationBound(eps, f)
    g._scalar = True
    return g.find_truncation_bound()


_BOUNDS_MEMO: dict = {}


def truncation_bounds(kernel_obj, eps, dtype=torch.float64, trunc_eps=None):
    """(Ltime, Lfreq): the eps-supports of k(r) and of |r|^(d-1) S(r)/S(0) (utils/kernels.py:91-104).  They do
    not depend on the domain size L, so a caller can run the two bisections while the device still reduces
    the bounding box of the points, and hand the result to get_xis(..., bounds=...)."""
    dim = kernel_obj.dimension
    if trunc_eps is None:
        trunc_eps = eps
    if hasattr(kernel_obj, "_scalar_closures"):
        # the two bisections are ~120 Python-level evaluations of the kernel (~60 us): a pure function of the kernel's
        # closed-form parameters, remembered for repeated fits with the same hyper-parameters (the shard-size step of a
        # multi-GPU fit is short enough for this host time to show)
        key = None
        if hasattr(kernel_obj, "_spectral_family"):
            try:
                key = (type(kernel_obj).__name__, dim, float(eps), float(trunc_eps),
                       tuple(float(v) for v in kernel_obj._spectral_family()))
            except Exception:
                key = None
        if key is not None and key in _BOUNDS_MEMO:
            return _BOUNDS_MEMO[key]
        kf, sf = kernel_obj._scalar_closures()
        Ltime = _scalar_bound(eps, kf)
        s0 = sf(0.0)
        Lfreq = _scalar_bound(trunc_eps, lambda r: abs(r ** (dim - 1)) * sf(r) / s0)
        if key is not None:
            if len(_BOUNDS_MEMO) >= 64:
                _BOUNDS_MEMO.clear()
            _BOUNDS_MEMO[key] = (Ltime, Lfreq)
    else:  # any object with tensor-valued kernel / spectral_density, exactly as the reference does it
        Ltime = GetTruncationBound(eps, kernel_obj.kernel, dtype=dtype).find_truncation_bound()
        zero = torch.tensor(0, device="cpu", dtype=dtype)
        Lfreq = GetTruncationBound(
            trunc_eps, lambda r: abs(r ** (dim - 1)) * kernel_obj.spectral_density(r)
            / kernel_obj.spectral_density(zero), dtype=dtype).find_truncation_bound()
    return Ltime, Lfreq


def get_xis(kernel_obj, eps, L, use_integral=False, l2scaled=False, dtype=torch.float64, trunc_eps=None,
            bounds=None):
    """1-D equispaced Fourier nodes for tolerance eps on a domain of size L (utils/kernels.py:72-143)."""
    dim = kernel_obj.dimension
    if trunc_eps is None:
        trunc_eps = eps
    if isinstance(L, torch.Tensor):
        L_val = L
    else:
        L_val = float(L)
    if use_integral:
        Ltime, Lfreq = bounds if bounds is not None else truncation_bounds(kernel_obj, eps, dtype, trunc_eps)
        h_spacing = 1 / (L_val + Ltime)
        hm = math.ceil(Lfreq / h_spacing)
    else:
        from kernels.matern import Matern
        from kernels.squared_exponential import SquaredExponential
        if isinstance(kernel_obj, Matern):      # heuristics of utils/kernels.py:106-121
            l, nu = kernel_obj.get_hyper("lengthscale"), kernel_obj.nu
            eps_use = eps / kernel_obj.get_hyper("variance")
            if l2scaled:
                s0 = kernel_obj._spectral_density_scalar(0.0)
                rl2sq = ((2 * nu / math.pi / l ** 2) ** (dim / 2) * s0 ** 2 / 2
                         * math.gamma(dim / 2 + 2 * nu) / math.gamma(dim + 2 * nu) * 2 ** (-dim / 2))
                eps_use = eps * math.sqrt(rl2sq)
            h_spacing = 1 / (L_val + 0.85 * l / math.sqrt(nu) * math.log(1 / eps_use))
            hm = math.ceil((math.pi ** (nu + dim / 2) * l ** (2 * nu) * eps_use / 0.15) ** (-1 / (2 * nu + dim / 2))
                           / h_spacing)
        elif isinstance(kernel_obj, SquaredExponential):   # utils/kernels.py:123-134
            l = kernel_obj.get_hyper("lengthscale")
            eps_use = eps / kernel_obj.get_hyper("variance")
            if l2scaled:
                rl2sq = kernel_obj._kernel_scalar(0.0) ** 2 * (math.sqrt(math.pi) * l ** 2) ** dim
                eps_use = eps * math.sqrt(rl2sq)
            h_spacing = 1 / (L_val + l * math.sqrt(2 * math.log(4 * dim * 3 ** dim / eps_use)))
            hm = math.ceil(math.sqrt(math.log(dim * (4 ** (dim + 1)) / eps_use) / 2) / math.pi / l / h_spacing)
        else:
            raise ValueError("heuristic grid is only defined for Matern and SquaredExponential kernels")
    xis = torch.arange(-hm, hm + 1, device="cpu", dtype=dtype) * h_spacing
    return xis, h_spacing, xis.numel()
