"""Pin the CPU oracle (oracle/efgp_oracle.py) against the golden vectors produced by the unmodified
reference (tests/golden/make_golden.py).  CPU only."""
import os

import numpy as np
import pytest
import torch

from conftest import oracle, GOLDEN

O = oracle()
NAMES = ["se", "matern12", "matern32", "matern52"]


def rel(a, b):
    a, b = torch.as_tensor(a), torch.as_tensor(b)
    return float(torch.linalg.norm((a - b).reshape(-1)) / (torch.linalg.norm(b.reshape(-1)) + 1e-300))


def test_get_xis_matches_reference():
    rows = np.load(os.path.join(GOLDEN, "host.npz"))["get_xis"]
    for row in rows[::3]:
        ki, d, ell, eps, L, h, mtot, x0, x1 = row
        k = O.OracleKernel(NAMES[int(ki)], int(d), ell, 1.3, param_dtype=torch.float32)
        xis, hh, mt = O.get_xis(k, eps, L)
        assert mt == int(mtot)
        assert abs(hh - h) <= 1e-15 * abs(h)
        assert abs(float(xis[0]) - x0) <= 1e-12 * abs(x0)


def test_spectral_density_and_grad():
    g = np.load(os.path.join(GOLDEN, "host.npz"))
    for name in NAMES:
        for d in (1, 2, 3):
            k = O.OracleKernel(name, d, 0.17, 1.7, param_dtype=torch.float32)
            t = torch.linspace(-3.0, 3.0, 7, dtype=torch.float64)
            xis = torch.stack(torch.meshgrid(*(t for _ in range(d)), indexing="ij"), -1).view(-1, d)
            assert rel(k.spectral_density(xis), g[f"sd_{name}_{d}"]) < 1e-14
            assert rel(k.spectral_grad(xis), g[f"sg_{name}_{d}"]) < 1e-13
            assert rel(k.kernel(torch.linspace(0, 2, 9, dtype=torch.float64)), g[f"k_{name}_{d}"]) < 1e-14


def test_toeplitz_matches_reference():
    g = np.load(os.path.join(GOLDEN, "toeplitz_cg.npz"))
    for tag in ("1d", "2d", "3d", "2d_rect"):
        T = O.OracleToeplitz(torch.from_numpy(g[f"toep_v_{tag}"]))
        assert T.fft_shape == list(g[f"toep_fftshape_{tag}"])
        y = T(torch.from_numpy(g[f"toep_x_{tag}"]))
        assert rel(y, g[f"toep_y_{tag}"]) < 1e-13


def test_cg_matches_reference():
    g = np.load(os.path.join(GOLDEN, "toeplitz_cg.npz"))
    A = torch.from_numpy(g["cg_A"])
    b1 = torch.from_numpy(g["cg_b1"])
    bB = torch.from_numpy(g["cg_bB"])
    dg = torch.diagonal(A).real
    for tol in (1e-3, 1e-8):
        for pre in (0, 1):
            Minv = (lambda v: v / dg) if pre else None
            x, it = O.cg_solve(lambda v: A @ v, b1, torch.zeros_like(b1), tol=tol, M_inv=Minv)
            assert it == int(g[f"cg_it1_{tol}_{pre}"])
            assert rel(x, g[f"cg_x1_{tol}_{pre}"]) < 1e-12
            x, it = O.cg_solve(lambda V: V @ A.T, bB, torch.zeros_like(bB), tol=tol, M_inv=Minv)
            assert it == int(g[f"cg_itB_{tol}_{pre}"])
            assert rel(x, g[f"cg_xB_{tol}_{pre}"]) < 1e-12


@pytest.mark.parametrize("case", ["cfg1_se1d", "se2d", "usa_f32", "matern52_3d"])
def test_model_matches_reference(case):
    g = np.load(os.path.join(GOLDEN, f"model_{case}.npz"))
    dt = torch.float32 if "float32" in str(g["dtype"]) else torch.float64
    x = torch.from_numpy(g["x"]).to(dt)
    y = torch.from_numpy(g["y"]).to(dt)
    x_new = torch.from_numpy(g["x_new"]).to(dt)
    d = x.shape[1]
    k = O.OracleKernel(str(g["kernel"]), d, float(g["ell"]), float(g["var"]), param_dtype=dt)
    sig2 = float(torch.exp(torch.log(torch.tensor(float(g["sig2"]), dtype=dt))))
    f64 = dt == torch.float64
    tol = 1e-10 if f64 else 1e-5
    neps = 1e-12 if f64 else 1e-6
    f = O.fit(x, y, k, sig2, float(g["eps"]), neps, cg_tol=tol)
    assert f.mtot == int(g["mtot"])
    assert rel(f.ws, g["ws"]) < (1e-13 if f64 else 1e-6)
    assert rel(f.beta, g["beta"]) < (1e-8 if f64 else 1e-3)
    mean = O.predict_mean(f, x_new, neps)
    assert rel(mean, g["mean"]) < (1e-9 if f64 else 1e-4)
    torch.manual_seed(321)
    vs = O.variance_stochastic(f, x_new, 24, tol, 2000, neps)
    assert rel(vs, g["var_stochastic"]) < (1e-7 if f64 else 1e-3)
    vr = O.variance_regular(f, x_new[:32], tol, 2000)
    assert rel(vr, g["var_regular"]) < (1e-7 if f64 else 1e-3)
    torch.manual_seed(1234)
    st = {}
    sig2_t = torch.exp(torch.log(torch.tensor(float(g["sig2"]), dtype=dt)))
    gr = O.gradient(x, y, k, sig2_t, float(g["eps"]), 4, neps, tol, stats=st)
    pos = torch.tensor([k.lengthscale, k.variance, float(sig2_t)], dtype=dt)
    raw = gr * pos
    assert rel(raw, g["grad_T4"]) < (1e-6 if f64 else 2e-3)
    if f64:
        assert abs(st["mean_cg_iters"] - int(g["grad_T4_iters"][0])) <= 1
        assert abs(st["trace_cg_iters"] - int(g["grad_T4_iters"][1])) <= 1


FAST_CASES = [(1, 3000, 35, 1e-4), (2, 5000, 17, 1e-4), (2, 5000, 16, 6e-8), (3, 2000, 9, 1e-6), (2, 3000, 21, 1e-12)]


@pytest.mark.parametrize("d,N,n,eps", FAST_CASES)
def test_fast_cpu_nufft_matches_exact_nudft(d, N, n, eps):
    """oracle/fast_nufft.py (C spreader + numpy FFT: the timed CPU baseline) against the exact NUDFT,
    within the same 10*eps relative l2 the GPU path is held to."""
    import fast_nufft as F
    from pytorch_finufft import functional as E
    g = torch.Generator().manual_seed(5)
    x = (torch.rand(d, N, dtype=torch.float64, generator=g) * 2 - 1) * 2.5
    c = torch.randn(N, dtype=torch.complex128, generator=g)
    f = torch.randn((2,) + (n,) * d, dtype=torch.complex128, generator=g)
    for modeord in (False, True):
        a = F.finufft_type1(x, c, (n,) * d, eps=eps, isign=-1, modeord=modeord)
        b = E.finufft_type1(x, c, (n,) * d, isign=-1, modeord=modeord)
        assert rel(a, b) < 10 * eps
        a2 = F.finufft_type2(x, f, eps=eps, isign=+1, modeord=modeord)
        b2 = E.finufft_type2(x, f, isign=+1, modeord=modeord)
        assert a2.shape == b2.shape and rel(a2, b2) < 10 * eps


def test_oracle_fit_with_fast_backend_matches_exact_backend():
    """The CPU-baseline configuration (reference algebra + fast CPU NUFFT) reproduces the exact-NUDFT fit."""
    import fast_nufft as F
    g = torch.Generator().manual_seed(11)
    x = torch.rand(3000, 2, dtype=torch.float64, generator=g) * 2 - 1
    y = torch.sin(3 * x[:, 0]) * torch.cos(4 * x[:, 1]) + 0.1 * torch.randn(3000, dtype=torch.float64, generator=g)
    k = O.OracleKernel("se", 2, 0.3, 1.0)
    fe = O.fit(x, y, k, 0.05, 1e-4, 1e-12, cg_tol=1e-10)
    ff = O.fit(x, y, k, 0.05, 1e-4, 1e-12, cg_tol=1e-10, backend=F)
    assert rel(ff.beta, fe.beta) < 1e-6
    assert rel(O.predict_mean(ff, x[:200], 1e-12, backend=F), O.predict_mean(fe, x[:200], 1e-12)) < 1e-7
