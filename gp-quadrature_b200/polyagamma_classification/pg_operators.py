"""Feature-space operators of the Polya-Gamma EFGP classifier/regressor on the B200 hot path.

Mirror of the operator builders of the reference's polyagamma_classification/pg_classifier.py -- the second
caller of the NUFFT / Toeplitz / CG primitives (pg_classifier.py:24-27).  Same names, arguments and results:

    _build_spectral_state        pg_classifier.py:310-374   grid, weights, F^*F generator, NUFFT closures
    _build_weighted_toeplitz     pg_classifier.py:377-384   generator of F^* diag(delta) F (one real type-1)
    _make_sigma_apply            pg_classifier.py:387-436   z -> Re F D (I + D F^* Delta F D)^-1 D F^* z
    _make_feature_space_solver   pg_classifier.py:439-504   (I + Ds F^* Omega F Ds) y = Ds q,  beta = y / Ds
    _predictive_mean             pg_classifier.py:653-668   Re F (ws2 * beta) at new points

What runs where: every system is solved by the library's lockstep batched CG (efgp_cg_solve) on the fused
operator I + ws * T_w(ws * .), which is operator mode 3 (A_var) with sigma^2 = 1 on the WEIGHTED Toeplitz plan.
The reference's default (use_exact_weighted_toeplitz_operator=False) applies the same matrix through two batched
NUFFTs per CG iteration; the two are the same linear operator (F^* Delta F is exactly Toeplitz on the grid), so
this module always builds the weighted generator -- one extra type-1 per E-step instead of 2 x iterations
NUFFTs -- and accepts the flag only for signature compatibility.  The E/M-step drivers, likelihood classes and
sklearn estimators of pg_classifier.py are host orchestration above these operators and are not part of the path.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable

import torch

from cg import ConjugateGradients
from efgpnd import NUFFT, ToeplitzND, compute_convolution_vector_vectorized_dD, _FusedOperator, _hermitian, _zero_guess
from efgp_b200 import ops as _ops
from utils.kernels import get_xis


@dataclass
class _SpectralState:
    """pg_classifier.py:61-76."""
    xis: torch.Tensor
    h: float
    mtot: int
    ws: torch.Tensor
    ws2: torch.Tensor
    Dprime: torch.Tensor
    toeplitz: ToeplitzND
    nufft_op: NUFFT
    fadj: Callable[[torch.Tensor], torch.Tensor]
    fwd: Callable[[torch.Tensor], torch.Tensor]
    fadj_batched: Callable[[torch.Tensor], torch.Tensor]
    fwd_batched: Callable[[torch.Tensor], torch.Tensor]
    out_shape: tuple


def _build_spectral_state(X, kernel, *, spectral_eps, trunc_eps, nufft_eps, rdtype, cdtype, device) -> _SpectralState:
    X = X.to(device=device, dtype=rdtype)
    if X.ndim == 1:
        X = X[:, None]
    d = X.shape[1]
    L = float((X.max(dim=0).values - X.min(dim=0).values).max())
    xis_1d, h, mtot = get_xis(kernel_obj=kernel, eps=spectral_eps, L=L, use_integral=True, l2scaled=False,
                              trunc_eps=trunc_eps)
    hm = (mtot - 1) // 2
    xis_1d = (torch.arange(-hm, hm + 1, device=device, dtype=torch.float64) * h).to(rdtype)
    xis = torch.stack(torch.meshgrid(*(xis_1d for _ in range(d)), indexing="ij"), dim=-1).view(-1, d)
    ws2_r = kernel.spectral_density(xis).to(dtype=rdtype) * h ** d
    ws_r = torch.sqrt(ws2_r)
    ws2, ws = ws2_r.to(cdtype), ws_r.to(cdtype)
    ws._efgp_real, ws._efgp_real_part = True, ws_r.contiguous()
    v_kernel = compute_convolution_vector_vectorized_dD(hm, X, h).to(dtype=cdtype)
    toeplitz = ToeplitzND(v_kernel, force_pow2=True)
    Dprime = (h ** d * kernel.spectral_grad(xis)).to(device=device, dtype=cdtype)
    out_shape = (mtot,) * d
    nufft_op = NUFFT(X, torch.zeros(d, device=device, dtype=rdtype), h, nufft_eps, cdtype=cdtype, device=device)
    M = ws.numel()
    # the library transforms are batched natively: the *_batched closures are the same call on (B, .) blocks
    fadj = lambda v: nufft_op.type1(v, out_shape=out_shape).reshape(-1)
    fwd = lambda fk: nufft_op.type2(fk, out_shape=out_shape)
    fadj_batched = lambda v: nufft_op.type1(v, out_shape=out_shape).reshape(v.shape[0], M)
    fwd_batched = lambda fk: nufft_op.type2(fk.reshape(-1, M), out_shape=out_shape)
    return _SpectralState(xis=xis, h=h, mtot=mtot, ws=ws, ws2=ws2, Dprime=Dprime, toeplitz=toeplitz,
                          nufft_op=nufft_op, fadj=fadj, fwd=fwd, fadj_batched=fadj_batched, fwd_batched=fwd_batched,
                          out_shape=out_shape)


def _build_weighted_toeplitz(delta: torch.Tensor, spectral: _SpectralState) -> ToeplitzND:
    """Generator of F^* diag(delta) F on (2 mtot - 1)^d: a type-1 with the (real) weights as strengths."""
    w = delta.flatten()
    w = w.real if w.is_complex() else w
    conv_shape = tuple(2 * n - 1 for n in spectral.out_shape)
    v_weighted = spectral.nufft_op.type1(w, out_shape=conv_shape)
    return ToeplitzND(v_weighted.to(dtype=spectral.ws.dtype), force_pow2=True)


def _identity_plus(ws_real, weighted_toeplitz, cdtype):
    """u -> u + ws * T_w(ws * u): operator mode A_var with sigma^2 = 1 on the weighted plan (one library call,
    and the form cg.ConjugateGradients hands to efgp_cg_solve)."""
    w = ws_real.to(cdtype)
    w._efgp_real, w._efgp_real_part = True, ws_real.contiguous()
    return _FusedOperator(w, weighted_toeplitz, 1.0, _ops.MODE_A_VAR, cdtype)


def _make_sigma_apply(spectral, delta, *, cg_tol, use_exact_weighted_toeplitz_operator=False, weighted_toeplitz=None):
    info = {"cg_iters": 0}
    cdtype = spectral.ws.dtype
    wt = weighted_toeplitz if weighted_toeplitz is not None else _build_weighted_toeplitz(delta, spectral)
    ws_r = spectral.ws.real.contiguous()
    A_feat = _identity_plus(ws_r, wt, cdtype)

    def sigma_apply(z: torch.Tensor) -> torch.Tensor:
        vector_input = z.dim() == 1
        if vector_input:
            z = z.unsqueeze(0)
        zr = z.real if z.is_complex() else z
        rhs = _hermitian(spectral.nufft_op.type1(zr, out_shape=spectral.out_shape, scale=ws_r).reshape(z.shape[0], -1))
        cg = ConjugateGradients(A_feat, rhs, x0=_zero_guess(rhs), tol=cg_tol, max_iter=2000, early_stopping=True)
        x = cg.solve()
        info["cg_iters"] = int(cg.iters_completed)
        result = spectral.nufft_op.type2(x, out_shape=spectral.out_shape, scale=ws_r, real_out=True)
        return result.squeeze(0) if vector_input else result

    return sigma_apply, info


def _make_feature_space_solver(delta, spectral, *, cg_tol, use_exact_weighted_toeplitz_operator=False,
                               weighted_toeplitz=None):
    cdtype = spectral.ws.dtype
    omega = delta.to(dtype=cdtype).flatten()
    D2_real = spectral.ws2.real
    eps_d = max(float(D2_real.mean()) * 1e-14, 1e-14)
    Ds_r = torch.sqrt(torch.clamp(D2_real, min=eps_d)).contiguous()
    Ds = Ds_r.to(cdtype)
    Dsinv = 1.0 / Ds
    info = {"cg_iters": 0}
    wt = weighted_toeplitz if weighted_toeplitz is not None else _build_weighted_toeplitz(delta, spectral)
    apply_IpS = _identity_plus(Ds_r, wt, cdtype)

    def apply_omega(v: torch.Tensor) -> torch.Tensor:
        return omega[:, None] * v if v.dim() == 2 else omega * v

    def solve_A_beta(q: torch.Tensor):
        rhs = Ds * q if q.dim() == 1 else q * Ds[None, :]
        cg = ConjugateGradients(apply_IpS, rhs, x0=_zero_guess(rhs), tol=cg_tol, max_iter=2000, early_stopping=True)
        y = cg.solve()
        info["cg_iters"] = int(cg.iters_completed)
        beta = Dsinv * y if q.dim() == 1 else y * Dsinv[None, :]
        return beta, cg.iters_completed

    return solve_A_beta, apply_omega, info


def _predictive_mean(X_new, beta_mean, spectral, *, nufft_eps):
    if X_new.ndim == 1:
        X_new = X_new[:, None]
    d = X_new.shape[1]
    eval_op = NUFFT(X_new, torch.zeros(d, device=X_new.device, dtype=X_new.dtype), spectral.h, nufft_eps,
                    cdtype=spectral.ws.dtype, device=X_new.device)
    return eval_op.type2(beta_mean, out_shape=spectral.out_shape, scale=spectral.ws2.real.contiguous(), real_out=True)
