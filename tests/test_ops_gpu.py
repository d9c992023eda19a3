"""GPU parity of the C-ABI operators (NUFFT type 1/2, Toeplitz matvec, batched CG) against the CPU
oracle on identical seeded inputs.  Tolerances are BASELINE.json's: NUFFT within 10*eps relative l2."""
import os

import numpy as np
import pytest
import torch

from conftest import oracle, GOLDEN

pytestmark = pytest.mark.gpu

O = oracle()
from pytorch_finufft import functional as nudft  # noqa: E402  (oracle stand-in, tests only)


def rel(a, b):
    a = torch.as_tensor(a).detach().cpu()
    b = torch.as_tensor(b).detach().cpu()
    return float(torch.linalg.norm((a - b).reshape(-1).to(torch.complex128)) /
                 (torch.linalg.norm(b.reshape(-1).to(torch.complex128)) + 1e-300))


@pytest.fixture(scope="module")
def ops():
    from efgp_b200 import ops
    return ops


def make_points(N, d, dt, seed, lo=-1.0, hi=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.rand(N, d, dtype=torch.float64, generator=g) * (hi - lo) + lo).to(dt)


CASES_T1 = [
    # d, N, n_modes, eps, dtype
    (1, 4000, 35, 1e-4, torch.float64),
    (1, 3000, 69, 6e-8, torch.float64),
    (2, 20000, 17, 1e-4, torch.float64),
    (2, 20000, 33, 6e-8, torch.float64),
    (2, 5000, 12, 1e-6, torch.float64),      # even mode count
    (2, 8000, 21, 1e-10, torch.float64),     # wide kernel -> direct path
    (2, 8000, 9, 1e-13, torch.float64),
    (3, 6000, 11, 1e-4, torch.float64),
    (3, 4000, 9, 1e-7, torch.float64),
    (2, 20000, 17, 1e-4, torch.float32),
    (3, 5000, 9, 1e-3, torch.float32),
    (1, 3000, 35, 1e-5, torch.float32),
]


@pytest.mark.parametrize("d,N,n,eps,dt", CASES_T1)
def test_type1_real_strengths(ops, d, N, n, eps, dt):
    h = 0.37
    x = make_points(N, d, dt, 1)
    g = torch.Generator().manual_seed(2)
    c = torch.randn(N, dtype=torch.float64, generator=g).to(dt)
    plan = ops.NufftPlan(x.cuda(), h)
    out = plan.type1(c.cuda(), (n,) * d, eps)
    phi = (2 * np.pi * h * x).T.contiguous()
    ref = nudft.finufft_type1(phi, c.to(ops.cmplx_of(dt)), (n,) * d, isign=-1, modeord=False)
    tol = 10 * max(eps, 2e-6 if dt == torch.float32 else 1e-14)
    assert out.shape == ref.shape
    assert rel(out, ref) < tol
    # all-ones strengths (Toeplitz generator path): centre entry is N
    ones = plan.type1(None, (n,) * d, eps)
    ref1 = nudft.finufft_type1(phi, torch.ones(N, dtype=ops.cmplx_of(dt)), (n,) * d, isign=-1, modeord=False)
    assert rel(ones, ref1) < tol


@pytest.mark.parametrize("d,N,n,eps,dt", CASES_T1)
def test_type2(ops, d, N, n, eps, dt):
    h = 0.41
    x = make_points(N, d, dt, 3)
    g = torch.Generator().manual_seed(4)
    cd = ops.cmplx_of(dt)
    f = torch.randn((n,) * d, dtype=torch.complex128, generator=g).to(cd)
    plan = ops.NufftPlan(x.cuda(), h)
    phi = (2 * np.pi * h * x).T.contiguous()
    ref = nudft.finufft_type2(phi, f, isign=+1, modeord=False)
    tol = 10 * max(eps, 2e-6 if dt == torch.float32 else 1e-14)
    out = plan.type2(f.cuda(), (n,) * d, eps)
    assert rel(out, ref) < tol
    out_r = plan.type2(f.cuda(), (n,) * d, eps, real_out=True)
    assert out_r.dtype == dt
    assert rel(out_r, ref.real) < tol * float(torch.linalg.norm(ref) / torch.linalg.norm(ref.real))


def test_type1_batched_complex_modeord_isign_scale(ops):
    d, N, n, eps = 2, 6000, (13, 18), 1e-6
    dt = torch.float64
    x = make_points(N, d, dt, 5, 0.0, 1.0)       # uncentred cloud: exercises the periodic fold
    g = torch.Generator().manual_seed(6)
    c = torch.randn(3, N, dtype=torch.complex128, generator=g)
    h = 0.9
    plan = ops.NufftPlan(x.cuda(), h)
    phi = (2 * np.pi * h * x).T.contiguous()
    for isign in (-1, 1):
        for modeord in (False, True):
            ref = nudft.finufft_type1(phi, c, n, isign=isign, modeord=modeord)
            out = plan.type1(c.cuda(), n, eps, isign=isign, modeord=modeord)
            assert out.shape == (3,) + n
            assert rel(out, ref) < 10 * eps
    sc = torch.rand(n[0] * n[1], dtype=dt, generator=g) + 0.5
    out = plan.type1(c.cuda(), n, eps, scale=sc.cuda())
    ref = nudft.finufft_type1(phi, c, n, isign=-1, modeord=False) * sc.view(n)
    assert rel(out, ref) < 10 * eps


def test_type2_batched_modeord_isign_scale(ops):
    d, N, n, eps = 2, 5000, (15, 10), 1e-6
    dt = torch.float64
    x = make_points(N, d, dt, 7, -0.3, 1.4)
    g = torch.Generator().manual_seed(8)
    f = torch.randn((4,) + n, dtype=torch.complex128, generator=g)
    h = 0.55
    plan = ops.NufftPlan(x.cuda(), h)
    phi = (2 * np.pi * h * x).T.contiguous()
    for isign in (-1, 1):
        for modeord in (False, True):
            ref = nudft.finufft_type2(phi, f, isign=isign, modeord=modeord)
            out = plan.type2(f.cuda(), n, eps, isign=isign, modeord=modeord)
            assert out.shape == (4, N)
            assert rel(out, ref) < 10 * eps
            out_r = plan.type2(f.cuda(), n, eps, isign=isign, modeord=modeord, real_out=True)
            assert rel(out_r, ref.real) < 20 * eps
    sc = torch.rand(n[0] * n[1], dtype=dt, generator=g) + 0.5
    out = plan.type2(f.reshape(4, -1).cuda(), n, eps, scale=sc.cuda())
    ref = nudft.finufft_type2(phi, f * sc.view(n), isign=+1, modeord=False)
    assert rel(out, ref) < 10 * eps


def test_shared_sort_between_generator_and_rhs(ops):
    """The F*y transform reuses the finer Toeplitz-generator grid (one sort for both)."""
    d, N, m = 2, 30000, 8
    dt = torch.float64
    x = make_points(N, d, dt, 9)
    y = torch.randn(N, dtype=dt, generator=torch.Generator().manual_seed(10))
    h = 0.45
    plan = ops.NufftPlan(x.cuda(), h)
    phi = (2 * np.pi * h * x).T.contiguous()
    v = plan.type1(None, (4 * m + 1,) * d, 6e-8)
    Fy = plan.type1(y.cuda(), (2 * m + 1,) * d, 1e-4)
    mean = plan.type2(Fy, (2 * m + 1,) * d, 1e-4, real_out=True)
    assert rel(v, nudft.finufft_type1(phi, torch.ones(N, dtype=torch.complex128), (4 * m + 1,) * d)) < 6e-7
    refFy = nudft.finufft_type1(phi, y.to(torch.complex128), (2 * m + 1,) * d)
    assert rel(Fy, refFy) < 1e-3
    assert rel(mean, nudft.finufft_type2(phi, refFy).real) < 1e-3


def test_empty_and_tiny_inputs(ops):
    dt = torch.float64
    plan = ops.NufftPlan(torch.zeros(0, 2, dtype=dt).cuda(), 0.5)
    out = plan.type1(torch.zeros(0, dtype=dt).cuda(), (5, 5), 1e-6)
    assert out.shape == (5, 5) and float(out.abs().max()) == 0.0
    x = torch.tensor([[0.25, -0.5]], dtype=dt)
    plan = ops.NufftPlan(x.cuda(), 0.5)
    out = plan.type1(torch.tensor([2.0], dtype=dt).cuda(), (5, 5), 1e-9)
    phi = (2 * np.pi * 0.5 * x).T.contiguous()
    ref = nudft.finufft_type1(phi, torch.tensor([2.0 + 0j], dtype=torch.complex128), (5, 5))
    assert rel(out, ref) < 1e-8


def test_linearity_of_point_sharding(ops):
    """Type-1 is linear in the points: spreading two halves separately and adding equals the whole
    (the identity behind the multi-GPU point sharding)."""
    d, N, n, eps = 2, 40000, 21, 1e-6
    dt = torch.float64
    x = make_points(N, d, dt, 11)
    c = torch.randn(N, dtype=dt, generator=torch.Generator().manual_seed(12))
    whole = ops.NufftPlan(x.cuda(), 0.4).type1(c.cuda(), (n,) * d, eps)
    a = ops.NufftPlan(x[: N // 3].cuda(), 0.4).type1(c[: N // 3].cuda(), (n,) * d, eps)
    b = ops.NufftPlan(x[N // 3:].cuda(), 0.4).type1(c[N // 3:].cuda(), (n,) * d, eps)
    assert rel(a + b, whole) < 1e-12


# ----------------------------------------------------------------------------- Toeplitz / CG
@pytest.mark.parametrize("tag", ["1d", "2d", "3d", "2d_rect"])
def test_toeplitz_matches_reference_golden(ops, tag):
    g = np.load(os.path.join(GOLDEN, "toeplitz_cg.npz"))
    v = torch.from_numpy(g[f"toep_v_{tag}"]).cuda()
    T = ops.ToeplitzPlan(v)
    assert T.fft_shape == list(g[f"toep_fftshape_{tag}"])
    y = T.apply(torch.from_numpy(g[f"toep_x_{tag}"]).cuda())
    assert rel(y, g[f"toep_y_{tag}"]) < 1e-13


@pytest.mark.parametrize("dt", [torch.float64, torch.float32])
@pytest.mark.parametrize("d,m", [(1, 17), (2, 8), (3, 3)])
def test_operator_modes_and_cg(ops, d, m, dt):
    cd = ops.cmplx_of(dt)
    N = 3000
    x = make_points(N, d, torch.float64, 13)
    h = 0.45
    phi = (2 * np.pi * h * x).T.contiguous()
    v = nudft.finufft_type1(phi, torch.ones(N, dtype=torch.complex128), (4 * m + 1,) * d).to(cd)
    n = 2 * m + 1
    M = n ** d
    g = torch.Generator().manual_seed(14)
    ws = (torch.rand(M, dtype=torch.float64, generator=g) * 0.05 + 0.01).to(dt)
    sig2 = 0.3
    Tref = O.OracleToeplitz(v)
    Gv, A_mean, A_var = O.make_ops(ws.to(cd), Tref, sig2)
    T = ops.ToeplitzPlan(v.cuda())
    X = torch.randn(5, M, dtype=torch.complex128, generator=g).to(cd)
    tol = 1e-12 if dt == torch.float64 else 2e-5
    assert rel(T.apply(X.cuda(), ops.MODE_T), Tref(X)) < tol
    assert rel(T.apply(X.cuda(), ops.MODE_GV, ws.cuda()), Gv(X)) < tol
    assert rel(T.apply(X.cuda(), ops.MODE_A_MEAN, ws.cuda(), sig2), A_mean(X)) < tol
    assert rel(T.apply(X.cuda(), ops.MODE_A_VAR, ws.cuda(), sig2), A_var(X)) < tol
    # CG: batched with Jacobi, per-row early stop
    c0 = v[tuple((s - 1) // 2 for s in v.shape)].real
    dg = (c0 * ws ** 2 + sig2)
    Minv = lambda r: r / dg
    B = torch.randn(6, M, dtype=torch.complex128, generator=g).to(cd)
    B[2] *= 1e-3
    B[4] = 0
    cg_tol = 1e-8 if dt == torch.float64 else 1e-4
    xr, itr = O.cg_solve(A_mean, B, torch.zeros_like(B), tol=cg_tol, M_inv=Minv)
    xg, itg = T.cg(B.cuda(), torch.zeros_like(B).cuda(), ops.MODE_A_MEAN, ws.cuda(), sig2, cg_tol,
                   minv=(1.0 / dg).cuda())
    assert abs(itg - itr) <= (0 if dt == torch.float64 else 2)
    assert rel(xg, xr) < (1e-7 if dt == torch.float64 else 1e-2)
    # single right-hand side, warm start, no preconditioner, A_var
    b1 = B[0]
    x0 = 0.5 * xr[0]
    xr1, itr1 = O.cg_solve(A_var, b1, x0, tol=cg_tol)
    xg1, itg1 = T.cg(b1.cuda(), x0.cuda(), ops.MODE_A_VAR, ws.cuda(), sig2, cg_tol)
    assert abs(itg1 - itr1) <= (0 if dt == torch.float64 else 2)
    assert rel(xg1, xr1) < (1e-7 if dt == torch.float64 else 1e-2)
    # max_iter cap and early_stopping=False
    xr2, itr2 = O.cg_solve(A_mean, B, torch.zeros_like(B), tol=1e-30, max_iter=7, M_inv=Minv)
    xg2, itg2 = T.cg(B.cuda(), torch.zeros_like(B).cuda(), ops.MODE_A_MEAN, ws.cuda(), sig2, 1e-30, max_iter=7,
                     minv=(1.0 / dg).cuda())
    assert itg2 == itr2 == 7
    assert rel(xg2, xr2) < (1e-9 if dt == torch.float64 else 1e-3)


def test_points_outside_one_period_and_clustered(ops):
    """phi may exceed one period (x_new outside the training box) and real data are
    clustered (many points per cell in a few cells, empty cells elsewhere)."""
    g = torch.Generator().manual_seed(9)
    N, n, eps, h = 30000, 19, 1e-6, 0.45
    centres = torch.rand(6, 2, dtype=torch.float64, generator=g) * 2 - 1
    x = centres[torch.randint(0, 6, (N,), generator=g)] + 0.01 * torch.randn(N, 2, dtype=torch.float64, generator=g)
    x[: N // 10] = torch.rand(N // 10, 2, dtype=torch.float64, generator=g) * 9 - 4.5     # several periods away
    c = torch.randn(N, dtype=torch.float64, generator=g)
    plan = ops.NufftPlan(x.cuda(), h)
    phi = (2 * np.pi * h * x).T.contiguous()
    ref = nudft.finufft_type1(phi, c.to(torch.complex128), (n, n), isign=-1, modeord=False)
    assert rel(plan.type1(c.cuda(), (n, n), eps), ref) < 10 * eps
    f = torch.randn(n, n, dtype=torch.complex128, generator=g)
    ref2 = nudft.finufft_type2(phi, f, isign=+1, modeord=False)
    assert rel(plan.type2(f.cuda(), (n, n), eps), ref2) < 10 * eps


@pytest.mark.parametrize("d,N", [(2, 5_000_000), (3, 2_000_000), (1, 3_000_000)])
def test_large_n_size_independent_properties(ops, d, N):
    """At sizes the exact oracle cannot reach: (i) the generator's centre entry is N and F*1 is Hermitian,
    (ii) type-2 of a single mode is that plane wave at every point (exercises the whole sort / permutation
    machinery: a misplaced point shows up as a wrong phase), (iii) adjointness <F* c, f> = <c, F f>."""
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.rand(N, d, dtype=torch.float64, device="cuda", generator=g) * 2 - 1
    h, eps, n = 0.43, 1e-7, 15
    plan = ops.NufftPlan(x, h)
    v = plan.type1(None, (n,) * d, eps)
    centre = tuple([n // 2] * d)
    assert abs(v[centre].real.item() - N) < 10 * eps * N and abs(v[centre].imag.item()) < 10 * eps * N
    assert rel(v, torch.conj(torch.flip(v, dims=tuple(range(d))))) < 10 * eps
    k = [3, -2, 1][:d]
    f = torch.zeros((n,) * d, dtype=torch.complex128, device="cuda")
    f[tuple(n // 2 + kk for kk in k)] = 1.0
    wave = torch.exp(1j * 2 * np.pi * h * (x * torch.tensor(k, dtype=torch.float64, device="cuda")).sum(1))
    assert rel(plan.type2(f, (n,) * d, eps), wave) < 10 * eps
    c = torch.randn(N, dtype=torch.float64, device="cuda", generator=g)
    fr = torch.randn((n,) * d, dtype=torch.complex128, device="cuda", generator=g)
    lhs = torch.vdot(fr.reshape(-1), plan.type1(c, (n,) * d, eps).reshape(-1))          # <f, F* c>
    rhs = torch.vdot(plan.type2(fr, (n,) * d, eps), c.to(torch.complex128))              # <F f, c>
    assert abs(lhs - rhs) / abs(lhs) < 20 * eps
