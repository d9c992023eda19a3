"""Large-N GP prior sampler through the EFGP spectral approximation (one batched type-2 NUFFT on the device).

Mirror of sample_gp_spectral_approx in the reference's vanilla_gp_sampling.py:166-260.
The dense O(N^3) samplers of that file (sample_gp_fast, sample_gp_matern) are toy baselines and are not part of the path.
"""
import math
from typing import Callable, Optional

import torch

import efgpnd as _efgpnd
from efgpnd import NUFFT
from kernels import SquaredExponential
from utils.kernels import get_xis


def mean_func_zero(x: torch.Tensor) -> torch.Tensor:
    """Zero mean (vanilla_gp_sampling.py:11-21)."""
    return torch.zeros(x.shape[0], device=x.device, dtype=x.dtype)


def sample_gp_spectral_approx(x: torch.Tensor, mean_func: Callable[[torch.Tensor], torch.Tensor] = mean_func_zero,
                              num_samples: int = 1, length_scale: float = 1.0, variance: float = 1.0,
                              spectral_eps: float = 1e-4, trunc_eps: float = 1e-4, nufft_eps: float = 1e-7,
                              seed: Optional[int] = None) -> torch.Tensor:
    """Draws from K ~ F D^2 F^*: sqrt(2) Re F (ws * c), c proper complex Gaussian.  Returns (n, num_samples), or (n,)
    for one sample.  The coefficients are drawn with the reference's calls (two torch.randn blocks, :249-251) on
    efgpnd.PROBE_DEVICE if set, else on x.device."""
    if num_samples <= 0:
        raise ValueError("num_samples must be positive.")
    if x.ndim == 1:
        x = x.unsqueeze(-1)
    if seed is not None:
        torch.manual_seed(seed)
    device, rdtype = x.device, x.dtype
    n_points, dimension = x.shape
    mean = mean_func(x).to(device=device, dtype=rdtype)
    L = float((x.max(dim=0).values - x.min(dim=0).values).max())
    if L <= 1e-9:
        L = 1.0
    kernel = SquaredExponential(dimension=dimension, init_lengthscale=float(length_scale), init_variance=float(variance))
    xis_1d, h, mtot = get_xis(kernel_obj=kernel, eps=spectral_eps, L=L, use_integral=True, l2scaled=False,
                              trunc_eps=trunc_eps)
    hm = (mtot - 1) // 2
    xis_1d = (torch.arange(-hm, hm + 1, device=device, dtype=torch.float64) * h).to(rdtype)
    xis = torch.stack(torch.meshgrid(*(xis_1d for _ in range(dimension)), indexing="ij"), dim=-1).view(-1, dimension)
    ws_r = torch.sqrt(kernel.spectral_density(xis).to(dtype=rdtype) * (h ** dimension)).contiguous()
    out_shape = (mtot,) * dimension
    nufft = NUFFT(x, torch.zeros(dimension, device=device, dtype=rdtype), h, nufft_eps, device=device)
    draw_dev = torch.device(_efgpnd.PROBE_DEVICE) if _efgpnd.PROBE_DEVICE is not None else device
    coeff_shape = (num_samples, ws_r.numel())
    coeff_real = torch.randn(coeff_shape, device=draw_dev, dtype=rdtype)
    coeff_imag = torch.randn(coeff_shape, device=draw_dev, dtype=rdtype)
    coeffs = torch.complex(coeff_real, coeff_imag).to(device) / math.sqrt(2.0)
    latent = nufft.type2(coeffs, out_shape=out_shape, scale=ws_r, real_out=True).reshape(num_samples, n_points)
    samples = mean.unsqueeze(0) + math.sqrt(2.0) * latent
    if num_samples == 1:
        return samples.squeeze(0)
    return samples.T.contiguous()
