"""TEST INFRASTRUCTURE ONLY -- CPU restatement (torch, float64/float32 on the host) of the
reference's EFGP hot path.  Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline /
reference arm may import this module; the product never does.

Pinned against the golden fixtures in tests/golden/ (generated from the UNMODIFIED reference by
tests/golden/make_golden.py) by tests/test_oracle.py.

Every function cites the reference lines it restates (paths relative to the reference tree).
The NUFFTs default to the exact nonuniform DFT of oracle/pytorch_finufft (FINUFFT itself is an
un-vendored dependency: finufft>=1.2.0 / pytorch-finufft, setup.py:17-18); `nufft=` lets the
bench substitute the fast C spreader in oracle/cnufft.c.
"""
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from pytorch_finufft import functional as exact_nudft  # noqa: E402


def cmplx(rd):
    return torch.complex64 if rd == torch.float32 else torch.complex128


# ----------------------------------------------------------------------------- kernels
class OracleKernel:
    """SE / Matern spectral densities (kernels/squared_exponential.py:64-123, kernels/matern.py:55-168)."""

    def __init__(self, name, dimension, lengthscale, variance, param_dtype=torch.float64):
        self.name = name
        self.dimension = dimension
        self.nu = {"se": None, "matern12": 0.5, "matern32": 1.5, "matern52": 2.5}[name]
        # hypers live as logs in a GPParams vector of the model dtype (kernel_params.py:39, efgpnd.py:447-449)
        # and are read back as exp(raw) (kernel.py:113-115)
        rt = lambda v: float(torch.exp(torch.log(torch.tensor(float(v), dtype=param_dtype))).item())
        self.lengthscale = rt(lengthscale)
        self.variance = rt(variance)

    def kernel(self, r):
        ell, var = self.lengthscale, self.variance
        if self.nu is None:
            return var * torch.exp(-0.5 * (r / ell) ** 2)
        s = torch.abs(r) / ell
        if self.nu == 0.5:
            return var * torch.exp(-s)
        if self.nu == 1.5:
            return var * (1 + math.sqrt(3) * s) * torch.exp(-math.sqrt(3) * s)
        return var * (1 + math.sqrt(5) * s + 5 * s ** 2 / 3) * torch.exp(-math.sqrt(5) * s)

    def spectral_density(self, xi):
        if xi.ndim == 1:
            xi = xi[:, None]
        q = (xi ** 2).sum(-1)
        d, ell, var = self.dimension, self.lengthscale, self.variance
        if self.nu is None:
            return (2 * math.pi * ell ** 2) ** (d / 2) * var * torch.exp(-(2 * math.pi) ** 2 * ell ** 2 * q / 2)
        nu = self.nu
        c = ((2 * math.sqrt(math.pi)) ** d * math.gamma(nu + d / 2) * (2 * nu) ** nu
             / (math.gamma(nu) * ell ** (2 * nu)))
        return var * c * (2 * nu / ell ** 2 + 4 * math.pi ** 2 * q) ** (-(nu + d / 2))

    def spectral_grad(self, xi):
        if xi.ndim == 1:
            xi = xi[:, None]
        q = (xi ** 2).sum(-1)
        d, ell, var = self.dimension, self.lengthscale, self.variance
        S = self.spectral_density(xi)
        if self.nu is None:
            dl = S * (d / ell - (2 * math.pi) ** 2 * ell * q)
        else:
            nu = self.nu
            den = 2 * nu / ell ** 2 + 4 * math.pi ** 2 * q
            dl = S * (-2 * nu / ell + (nu + d / 2) * (4 * nu / ell ** 3) / den)
        return torch.stack([dl, S / var], -1)


def _bisect(f, eps, hi=1000.0, iters=200):
    """utils/kernels.py:27-69: double the bracket up to 10 times, then 200 bisection steps; returns
    the last midpoint."""
    lo = 0.0
    for _ in range(10):
        if f(hi) > eps:
            hi *= 2
        else:
            break
    mid = 0.5 * (lo + hi)
    for _ in range(iters):
        mid = 0.5 * (lo + hi)
        if f(mid) > eps:
            lo = mid
        else:
            hi = mid
    return mid


def get_xis(kern, eps, L):
    """utils/kernels.py:72-143 with use_integral=True: returns (xis_1d float64, h, mtot)."""
    t = lambda v: torch.tensor(v, dtype=torch.float64)
    Ltime = _bisect(lambda r: float(kern.kernel(t(r))), eps)
    h = 1.0 / (L + Ltime)
    d = kern.dimension
    s0 = float(kern.spectral_density(t(0.0).reshape(1))[0])
    Lfreq = _bisect(lambda r: abs(r ** (d - 1)) * float(kern.spectral_density(t(r).reshape(1))[0]) / s0, eps)
    hm = math.ceil(Lfreq / h)
    xis = torch.arange(-hm, hm + 1, dtype=torch.float64) * h
    return xis, h, 2 * hm + 1


# ----------------------------------------------------------------------------- Toeplitz
class OracleToeplitz:
    """efgpnd.py:1244-1393."""

    def __init__(self, v, force_pow2=True):
        self.Ls = list(v.shape)
        self.d = len(self.Ls)
        self.ns = [(L + 1) // 2 for L in self.Ls]
        self.size = math.prod(self.ns)
        self.fft_shape = [1 << (L - 1).bit_length() for L in self.Ls] if force_pow2 else list(self.Ls)
        self.dims = list(range(-self.d, 0))
        self.vhat = torch.fft.fftn(v, s=self.fft_shape, dim=self.dims)

    def __call__(self, x):
        flat = x.shape[-1] == self.size  # the reference tests the flat layout first (efgpnd.py:1345)
        lead = x.shape[:-1] if flat else x.shape[:-self.d]
        xb = x.reshape(*lead, *self.ns).to(self.vhat.dtype)
        y = torch.fft.ifftn(torch.fft.fftn(xb, s=self.fft_shape, dim=self.dims) * self.vhat, dim=self.dims)
        idx = [slice(None)] * len(lead) + [slice(n - 1, 2 * n - 1) for n in self.ns]
        y = y[tuple(idx)]
        return y.reshape(x.shape)


def make_ops(ws, toep, sigmasq):
    """efgpnd.py:1572-1631: Gv, A_mean, A_var and the Jacobi preconditioner factory."""
    Gv = lambda v: ws * toep(ws * v)
    A_mean = lambda v: Gv(v) + sigmasq * v
    A_var = lambda v: Gv(v) / sigmasq + v
    return Gv, A_mean, A_var


def jacobi(ws, sigmasq, c0):
    dg = c0 * ws.abs() ** 2 + sigmasq
    return lambda v: v / dg


# ----------------------------------------------------------------------------- CG
def cg_solve(A, b, x0, tol=1e-6, max_iter=None, early_stopping=True, M_inv=None):
    """cg.py:86-244.  Returns (x, iters_completed)."""
    tiny = 1e-16
    if b.dim() == 1:
        n = b.numel()
        max_iter = 2 * n if max_iter is None else max_iter
        x = x0.clone()
        r = b - A(x)
        z = M_inv(r) if M_inv else r.clone()
        p = z.clone()
        rz = torch.vdot(r, z).real
        bn = torch.linalg.norm(b).real
        den = bn if bn > 0 else torch.ones_like(bn)
        it = 0
        for it in range(max_iter):
            Ap = A(p)
            a = rz / (torch.vdot(p, Ap).real + tiny)
            x = x + a * p
            r = r - a * Ap
            if early_stopping and torch.linalg.norm(r).real / (den + tiny) < tol:
                break
            z = M_inv(r) if M_inv else r
            rz_new = torch.vdot(r, z).real
            p = z + (rz_new / (rz + tiny)) * p
            rz = rz_new
        return x, it + 1
    n = b.shape[1]
    max_iter = 2 * n if max_iter is None else max_iter
    x = x0.clone()
    r = b - A(x)
    z = M_inv(r) if M_inv else r.clone()
    p = z.clone()
    rz = (r.conj() * z).sum(1).real
    bn = torch.linalg.norm(b, dim=1).real
    den = torch.where(bn > 0, bn, torch.ones_like(bn))
    active = torch.ones(b.shape[0], dtype=torch.bool)
    it = 0
    for it in range(max_iter):
        idx = torch.where(active)[0]
        if idx.numel() == 0:
            break
        pa = p[idx]
        Ap = A(pa)
        a = rz[idx] / ((pa.conj() * Ap).sum(1).real + tiny)
        x[idx] += a[:, None] * pa
        r[idx] -= a[:, None] * Ap
        ra = r[idx]
        za = M_inv(ra) if M_inv else ra
        rz_new = (ra.conj() * za).sum(1).real
        p[idx] = za + (rz_new / (rz[idx] + tiny))[:, None] * pa
        rz[idx] = rz_new
        if early_stopping:
            rn = torch.linalg.norm(ra, dim=1).real
            done = (rn / (den[idx] + tiny) < tol) | (rn < 1e-12)
            active[idx[done]] = False
    return x, it + 1


# ----------------------------------------------------------------------------- model
class OracleNUFFT:
    """efgpnd.py:1423-1549 on top of a pluggable finufft-like backend."""

    def __init__(self, x, h, eps, backend=exact_nudft):
        self.phi = (2 * math.pi * h * x).T.contiguous()
        self.eps = eps
        self.cd = cmplx(x.dtype)
        self.be = backend

    def type1(self, vals, out_shape):
        return self.be.finufft_type1(self.phi, vals.to(self.cd), tuple(out_shape), eps=self.eps, isign=-1,
                                     modeord=False).to(self.cd)

    def type2(self, fk, out_shape):
        lead = fk.shape[:-1] if fk.ndim > 1 else ()
        return self.be.finufft_type2(self.phi, fk.to(self.cd).reshape(*lead, *out_shape), eps=self.eps, isign=+1,
                                     modeord=False)


class OracleFit:
    pass


def fit(x, y, kern, sigmasq, eps, nufft_eps, cg_tol=1e-4, backend=exact_nudft, precond=True, beta0=None):
    """EFGPND._compute_common_parameters, efgpnd.py:745-822."""
    rd = x.dtype
    cd = cmplx(rd)
    N, d = x.shape
    L = float((x.max(0).values - x.min(0).values).max())
    if L <= 1e-9:
        L = 1.0
    xis1, hf, mtot = get_xis(kern, eps, L)
    h = torch.tensor(hf, dtype=rd)
    xis1 = xis1.to(rd)
    xis = torch.stack(torch.meshgrid(*(xis1 for _ in range(d)), indexing="ij"), -1).view(-1, d)
    ws = torch.sqrt(kern.spectral_density(xis).to(cd) * h ** d).to(cd)
    OUT = (mtot,) * d
    op = OracleNUFFT(x, h, nufft_eps, backend)
    rhs = ws * op.type1(y, OUT).reshape(-1)
    m = (mtot - 1) // 2
    v = OracleNUFFT(x, h, 6e-8, backend).type1(torch.ones(N, dtype=cd), (4 * m + 1,) * d)
    toep = OracleToeplitz(v)
    _, A_mean, _ = make_ops(ws, toep, sigmasq)
    c0 = v[(tuple((s - 1) // 2 for s in v.shape))].real
    Minv = jacobi(ws, sigmasq, c0) if precond else None
    x0 = beta0.clone() if beta0 is not None and beta0.shape == rhs.shape else torch.zeros_like(rhs)
    beta, iters = cg_solve(A_mean, rhs, x0, tol=cg_tol, M_inv=Minv)
    f = OracleFit()
    f.beta, f.ws, f.xis, f.h, f.hf, f.mtot, f.toep, f.iters, f.v = beta, ws, xis, h, hf, mtot, toep, iters, v
    f.d, f.sigmasq, f.rd, f.cd = d, sigmasq, rd, cd
    f.rhs, f.A_mean = rhs, A_mean      # the assembled system, for residual checks of a foreign beta
    return f


def predict_mean(f, x_new, nufft_eps, backend=exact_nudft):
    """efgpnd.py:918-922."""
    op = OracleNUFFT(x_new, f.h, nufft_eps, backend)
    return op.type2(f.ws * f.beta, (f.mtot,) * f.d).real


def variance_stochastic(f, x_new, J, cg_tol, max_iter, nufft_eps, backend=exact_nudft):
    """diag_sums_nd + nufft_var_est_nd, efgpnd.py:1634-1679 (consumes the global torch RNG)."""
    _, _, A_var = make_ops(f.ws, f.toep, f.sigmasq)
    M = f.xis.shape[0]
    etas = (torch.randint(0, 2, (J, M)) * 2 - 1).to(f.xis.dtype)
    rhs = f.ws[None, :] * etas
    us, _ = cg_solve(A_var, rhs, torch.zeros_like(rhs), tol=cg_tol, max_iter=max_iter)
    gam = (f.ws[None, :] * us).view((J,) + (f.mtot,) * f.d)
    eta = etas.view((J,) + (f.mtot,) * f.d)
    s = (2 * f.mtot - 1,) * f.d
    dims = tuple(range(1, f.d + 1))
    R = torch.fft.ifftn(torch.fft.fftn(gam, s=s, dim=dims) * torch.conj(torch.fft.fftn(eta, s=s, dim=dims)),
                        s=s, dim=dims).mean(0)
    phi = (2 * math.pi * f.h * x_new).T.contiguous()
    return backend.finufft_type2(phi, R, eps=nufft_eps, isign=+1, modeord=True).real


def variance_regular(f, x_new, cg_tol, max_iter):
    """efgpnd.py:1805-1820."""
    _, _, A_var = make_ops(f.ws, f.toep, f.sigmasq)
    out = []
    for xb in torch.split(x_new, 8192, dim=0):
        fx = torch.exp(2j * math.pi * (xb @ f.xis.T)).to(f.cd)
        rhs = f.ws * fx.conj()
        g, _ = cg_solve(A_var, rhs, torch.zeros_like(rhs), tol=cg_tol, max_iter=max_iter)
        out.append(torch.real((fx * (f.ws * g)).sum(-1)).clamp_min(0.0))
    return torch.cat(out)


def gradient(x, y, kern, sigmasq, eps, T, nufft_eps, cg_tol, backend=exact_nudft, beta0=None, stats=None):
    """efgpnd_gradient_batched, efgpnd.py:61-264 (positive-space gradient of the negative log marginal;
    consumes the global torch RNG exactly like the reference: Z first, then V)."""
    rd = x.dtype
    cd = cmplx(rd)
    N, d = x.shape
    L = (x.max(0).values - x.min(0).values).max()
    xis1, h, mtot = get_xis(kern, eps, L)          # here h is a 0-d tensor of x's dtype (efgpnd.py:78-79, 95)
    xis = torch.stack(torch.meshgrid(*(xis1 for _ in range(d)), indexing="ij"), -1).view(-1, d)
    ws = torch.sqrt(kern.spectral_density(xis).to(cd) * h ** d)
    Dp = (h ** d * kern.spectral_grad(xis)).to(cd)
    OUT = (mtot,) * d
    op = OracleNUFFT(x, h, nufft_eps, backend)
    fadj = lambda v: op.type1(v, OUT).reshape(*v.shape[:-1], -1)
    fwd = lambda fk: op.type2(fk, OUT)
    m = (mtot - 1) // 2
    v = OracleNUFFT(x, h, 6e-8, backend).type1(torch.ones(N, dtype=cd), (4 * m + 1,) * d)
    toep = OracleToeplitz(v)
    _, A_mean, _ = make_ops(ws, toep, sigmasq)
    c0 = v[tuple((s - 1) // 2 for s in v.shape)].real
    Minv = jacobi(ws, sigmasq, c0)
    Fy = fadj(y.to(cd))
    rhs = ws * Fy
    x0 = beta0.clone() if beta0 is not None and beta0.shape == rhs.shape else torch.zeros_like(rhs)
    beta, it_mean = cg_solve(A_mean, rhs, x0, tol=cg_tol, M_inv=Minv)
    beta_raw = beta.clone()
    alpha = (y.to(cd) - fwd(ws * beta)) / sigmasq
    fa = (Fy - toep(ws * beta)) / sigmasq
    var = kern.variance
    t2 = torch.empty(3, dtype=rd)
    t2[0] = torch.vdot(fa, Dp[:, 0] * fa).real
    an = torch.vdot(alpha, alpha).real
    t2[1] = (torch.vdot(y.to(cd), alpha).real - sigmasq * an) / var
    t2[2] = an
    Z = torch.empty((T, N), dtype=rd).bernoulli_(0.5).mul_(2).sub_(1).to(cd)
    FZ = fadj(Z)
    DFZ = Dp[:, 0] * FZ
    rhs_k = fwd(DFZ)
    Bk = ws * toep(DFZ)
    V = torch.empty((T, ws.numel()), dtype=rd).bernoulli_(0.5).mul_(2).sub_(1).to(cd)
    Bn = ws * toep(ws * V)
    Ball = torch.cat([Bk, Bn], 0)
    Beta, it_tr = cg_solve(A_mean, Ball, torch.zeros_like(Ball), tol=cg_tol, M_inv=Minv)
    Bk_sol, Bn_sol = Beta[:T], Beta[T:]
    Alpha = (rhs_k - fwd(ws * Bk_sol)) / sigmasq
    t1 = torch.empty(3, dtype=rd)
    t1[0] = (Z * Alpha).sum(1).mean().real
    t1n = N / sigmasq - ((V.conj() * Bn_sol).sum(1).real / sigmasq).mean()
    t1[1] = (N - sigmasq * t1n) / var
    t1[2] = t1n
    if stats is not None:
        stats.update(mean_cg_iters=it_mean, trace_cg_iters=it_tr, trace_num_rhs=2 * T, mtot=mtot, mean_beta=beta_raw)
    return 0.5 * (t1 - t2)
