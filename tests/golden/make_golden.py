"""Generate the golden fixtures under tests/golden/ by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

The reference's NUFFT dependency (pytorch_finufft / FINUFFT) is not installable here, so the
exact-NUDFT stand-in in oracle/pytorch_finufft is put on sys.path under that name; everything
else (efgpnd.py, cg.py, kernels/, utils/kernels.py) is the reference's own code.  Outputs are
small .npz files committed beside this script; the GPU box never needs /root/reference.
"""
import io
import os
import sys
import contextlib

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, "/root/reference")

import efgpnd as ref  # noqa: E402
from cg import ConjugateGradients  # noqa: E402
from kernels.squared_exponential import SquaredExponential  # noqa: E402
from kernels.matern import Matern  # noqa: E402
from utils.kernels import get_xis  # noqa: E402

REFDATA = "/root/reference/data"


def npy(t):
    if isinstance(t, torch.Tensor):
        return t.detach().cpu().numpy()
    return np.asarray(t)


def make_kernel(name, d, ell, var):
    if name == "se":
        return SquaredExponential(dimension=d, init_lengthscale=ell, init_variance=var)
    nu = {"matern12": 0.5, "matern32": 1.5, "matern52": 2.5}[name]
    return Matern(dimension=d, init_lengthscale=ell, init_variance=var, nu=nu)


# --------------------------------------------------------------------------- host scalars
def golden_host():
    rows = []
    for name in ("se", "matern12", "matern32", "matern52"):
        for d in (1, 2, 3):
            for ell in (0.05, 0.1, 0.5):
                for eps in (1e-2, 1e-4, 1e-6):
                    for L in (1.0, 2.0):
                        if name == "matern12" and eps < 1e-2:
                            continue  # mtot explodes (algebraic decay); not a reference use case
                        k = make_kernel(name, d, ell, 1.3)
                        xis, h, mtot = get_xis(k, eps, L, use_integral=True, l2scaled=False)
                        rows.append((["se", "matern12", "matern32", "matern52"].index(name), d, ell, eps, L,
                                     float(h), int(mtot), float(xis[0]), float(xis[-1])))
    out = {"get_xis": np.array(rows, dtype=np.float64)}
    # spectral density / grad on a small 2-D and 3-D grid
    for name in ("se", "matern32", "matern52", "matern12"):
        for d in (1, 2, 3):
            k = make_kernel(name, d, 0.17, 1.7)
            g = torch.linspace(-3.0, 3.0, 7, dtype=torch.float64)
            xis = torch.stack(torch.meshgrid(*(g for _ in range(d)), indexing="ij"), -1).view(-1, d)
            out[f"sd_{name}_{d}"] = npy(k.spectral_density(xis))
            out[f"sg_{name}_{d}"] = npy(k.spectral_grad(xis))
            out[f"k_{name}_{d}"] = npy(k.kernel(torch.linspace(0, 2, 9, dtype=torch.float64)))
    # float32 round trip of hypers (kernel.py:137)
    k = make_kernel("se", 2, 0.5, 1.0)
    k.set_hyper("lengthscale", 0.1)
    out["hyper_roundtrip"] = np.array([k.get_hyper("lengthscale")])
    np.savez_compressed(os.path.join(HERE, "host.npz"), **out)


# --------------------------------------------------------------------------- Toeplitz + CG
def golden_toeplitz_cg():
    out = {}
    g = torch.Generator().manual_seed(7)
    for tag, Ls in (("1d", (9,)), ("2d", (13, 13)), ("3d", (9, 9, 9)), ("2d_rect", (9, 17))):
        v = torch.randn(*Ls, dtype=torch.complex128, generator=g)
        T = ref.ToeplitzND(v, force_pow2=True)
        x = torch.randn(3, T.size, dtype=torch.complex128, generator=g)
        out[f"toep_v_{tag}"] = npy(v)
        out[f"toep_x_{tag}"] = npy(x)
        out[f"toep_y_{tag}"] = npy(T(x))
        out[f"toep_fftshape_{tag}"] = np.array(T.fft_shape)
    # CG on a small Hermitian PD dense matrix, single and batched, with and without Jacobi
    n = 40
    B = torch.randn(n, n, dtype=torch.complex128, generator=g)
    A = B @ B.conj().T + 0.5 * torch.eye(n, dtype=torch.complex128)
    b1 = torch.randn(n, dtype=torch.complex128, generator=g)
    bB = torch.randn(6, n, dtype=torch.complex128, generator=g)
    bB[2] *= 1e-3
    out["cg_A"] = npy(A)
    out["cg_b1"] = npy(b1)
    out["cg_bB"] = npy(bB)
    diag = torch.diagonal(A).real
    Minv = lambda v: v / diag
    for tol in (1e-3, 1e-8):
        for pre in (0, 1):
            s = ConjugateGradients(lambda v: A @ v, b1, torch.zeros_like(b1), tol=tol,
                                   M_inv_apply=Minv if pre else None)
            out[f"cg_x1_{tol}_{pre}"] = npy(s.solve())
            out[f"cg_it1_{tol}_{pre}"] = np.array(s.iters_completed)
            s = ConjugateGradients(lambda V: V @ A.T, bB, torch.zeros_like(bB), tol=tol,
                                   M_inv_apply=Minv if pre else None)
            out[f"cg_xB_{tol}_{pre}"] = npy(s.solve())
            out[f"cg_itB_{tol}_{pre}"] = np.array(s.iters_completed)
    np.savez_compressed(os.path.join(HERE, "toeplitz_cg.npz"), **out)


# --------------------------------------------------------------------------- EFGPND end to end
def load_case(case):
    if case == "cfg1_se1d":  # BASELINE.json configs[0]; the file is 1-D
        d = torch.load(os.path.join(REFDATA, "gp_samples_5000_0.1_2_0.1.pt"))
        x, y = d[0].reshape(-1, 1), d[1]
        return x, y, "se", 0.1, 1.0, 0.1, 1e-4, torch.float64
    if case == "se2d":
        d = torch.load(os.path.join(REFDATA, "gp_samples_30000_0.5_2_0.2.pt"))
        x, y = d[0][:3000], d[1][:3000]
        return x, y, "se", 0.3, 1.2, 0.2, 1e-4, torch.float64
    if case == "matern32_2d":  # BASELINE.json configs[1] shape, sub-sampled
        d = torch.load(os.path.join(REFDATA, "gp_samples_100000_0.2_2_0.2.pt"))
        x, y = d[0][:4000], d[1][:4000]
        return x, y, "matern32", 0.4, 1.0, 0.2, 1e-3, torch.float64
    if case == "cfg2_full":  # BASELINE.json configs[1] at its REAL size: all 1e5 points, Matern-3/2, eps 1e-4 (mtot 101)
        d = torch.load(os.path.join(REFDATA, "gp_samples_100000_0.2_2_0.2.pt"))
        return d[0], d[1], "matern32", 0.2, 1.0, 0.2, 1e-4, torch.float64
    if case == "usa_f32":  # float32 mode on real data
        d = torch.load(os.path.join(REFDATA, "usa_temp_data.pt"))
        x, y = d["x"][:1500].float(), d["y"][:1500].float()
        y = (y - y.mean()) / y.std()
        x = (x - x.min(0).values) / (x.max(0).values - x.min(0).values).max()
        return x, y, "se", 0.15, 1.0, 0.3, 1e-3, torch.float32
    if case == "se3d":
        g = torch.Generator().manual_seed(11)
        x = torch.rand(1500, 3, dtype=torch.float64, generator=g) * 2 - 1
        y = torch.sin(3 * x[:, 0]) * torch.cos(2 * x[:, 1]) + 0.5 * x[:, 2] + 0.2 * torch.randn(1500, dtype=torch.float64, generator=g)
        return x, y, "se", 0.5, 1.0, 0.2, 1e-3, torch.float64
    if case == "matern52_3d":
        g = torch.Generator().manual_seed(12)
        x = torch.rand(1200, 3, dtype=torch.float64, generator=g) * 2 - 1
        y = torch.cos(2 * x[:, 0] + x[:, 1]) - 0.3 * x[:, 2] ** 2 + 0.1 * torch.randn(1200, dtype=torch.float64, generator=g)
        return x, y, "matern52", 0.9, 1.0, 0.1, 1e-2, torch.float64
    raise KeyError(case)


CASES = ["cfg1_se1d", "se2d", "matern32_2d", "usa_f32", "se3d", "matern52_3d", "cfg2_full"]


def golden_model(case):
    x, y, kname, ell, var, sig2, eps, dt = load_case(case)
    d = x.shape[1]
    out = {"x": npy(x), "y": npy(y), "kernel": np.array(kname), "ell": ell, "var": var, "sig2": sig2,
           "eps": eps, "dtype": np.array(str(dt))}
    k = make_kernel(kname, d, ell, var)
    # tight CG so the comparison is about the operators, not about where CG stopped
    opts = {"cg_tolerance": 1e-10 if dt == torch.float64 else 1e-5, "max_cg_iterations": 2000}
    nufft_eps = 1e-12 if dt == torch.float64 else 1e-6
    m = ref.EFGPND(x, y, k, sigmasq=sig2, eps=eps, nufft_eps=nufft_eps, opts=opts, estimate_params=False)
    g = torch.Generator().manual_seed(5)
    lo, hi = x.min(0).values, x.max(0).values
    x_new = (lo + (hi - lo) * torch.rand(257, d, dtype=torch.float64, generator=g).to(dt)).to(dt)
    out["x_new"] = npy(x_new)
    with contextlib.redirect_stdout(io.StringIO()):
        if d < 3:
            mean, _ = m.predict(x_new, return_variance=False)
        else:
            # efgpnd.py:908 computes mtot = int(M ** (1/d)), which truncates to mtot-1 for every odd
            # mtot > 3 when d == 3 (e.g. 1331 ** (1/3) = 10.999...), so the reference's predict()
            # raises in 3-D.  Drive the same reference pieces with the correct mtot instead.
            m._compute_common_parameters()
            mt = int(round(m._xis.shape[0] ** (1.0 / d)))
            hh = torch.tensor(m._xis.h_float, dtype=dt)
            op = ref.NUFFT(x_new, torch.zeros(d, dtype=dt), hh, nufft_eps)
            mean = op.type2(m._ws * m._beta, out_shape=(mt,) * d).real
        out["mean"] = npy(mean)
        out["beta"] = npy(m._beta)
        out["ws"] = npy(m._ws)
        out["mtot"] = np.array(int(round(m._xis.shape[0] ** (1.0 / d))))
        out["h"] = np.array(m._xis.h_float)
        cdt = ref._cmplx(dt)
        A_var = ref.create_A_var(m._ws, m._toeplitz, float(m._gp_params.sig2), cdt)
        hh = torch.tensor(m._xis.h_float, dtype=dt)
        common = dict(xis=m._xis, ws=m._ws, A_var=A_var, cg_tol=opts["cg_tolerance"],
                      max_cg_iter=opts["max_cg_iterations"], h=hh, xcen=torch.zeros(d, dtype=dt),
                      nufft_eps=nufft_eps, device=x.device, rdtype=dt, cdtype=cdt)
        torch.manual_seed(321)
        var_s = ref.compute_prediction_variance(x_new=x_new, variance_method="stochastic",
                                                hutchinson_probes=24, **common)
        out["var_stochastic"] = npy(var_s)
        var_r = ref.compute_prediction_variance(x_new=x_new[:32], variance_method="regular",
                                                hutchinson_probes=0, **common)
        out["var_regular"] = npy(var_r)
        # gradient with the verify_efgpnd_exact_small.py seeding convention (:254)
        for T in (4,):
            torch.manual_seed(1234)
            grads = m.compute_gradients(trace_samples=T, cg_tol=1e-10 if dt == torch.float64 else 1e-5,
                                        nufft_eps=nufft_eps, apply_gradients=False)
            out[f"grad_T{T}"] = npy(grads)
            st = m.last_gradient_stats
            out[f"grad_T{T}_iters"] = np.array([st["mean_cg_iters"], st["trace_cg_iters"], st["trace_num_rhs"]])
    # default-tolerance run (reference defaults: cg_tolerance 1e-4, nufft_eps 1e-4): records iteration counts
    k2 = make_kernel(kname, d, ell, var)
    m2 = ref.EFGPND(x, y, k2, sigmasq=sig2, eps=eps, nufft_eps=1e-4 if dt == torch.float64 else 1e-4,
                    estimate_params=False)
    with contextlib.redirect_stdout(io.StringIO()):
        if d < 3:
            mean2, _ = m2.predict(x_new, return_variance=False)
        else:
            m2._compute_common_parameters()
            mt = int(round(m2._xis.shape[0] ** (1.0 / d)))
            op = ref.NUFFT(x_new, torch.zeros(d, dtype=dt), torch.tensor(m2._xis.h_float, dtype=dt), 1e-4)
            mean2 = op.type2(m2._ws * m2._beta, out_shape=(mt,) * d).real
    out["mean_default"] = npy(mean2)
    np.savez_compressed(os.path.join(HERE, f"model_{case}.npz"), **out)
    print(case, "mtot", int(out["mtot"]), "grad", out["grad_T4"], "iters", out["grad_T4_iters"])


if __name__ == "__main__":
    torch.set_num_threads(8)
    only = sys.argv[1:]
    if not only:
        golden_host()
        golden_toeplitz_cg()
    for c in (only or CASES):
        golden_model(c)
