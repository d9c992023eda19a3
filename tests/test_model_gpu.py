"""End-to-end parity of the drop-in EFGPND (CUDA) against golden outputs of the UNMODIFIED reference
(tests/golden/make_golden.py).  Tolerances are BASELINE.json's: posterior mean 1e-6 relative l2 (fp64);
variance and gradient 1e-5 given the same probe seeds."""
import io
import os
import contextlib

import numpy as np
import pytest
import torch

from conftest import GOLDEN

pytestmark = pytest.mark.gpu

# cfg2_full: BASELINE.json configs[1] at its real size (all 1e5 points of gp_samples_100000_0.2_2_0.2.pt, Matern-3/2,
# eps 1e-4: mtot 101, 201^2 generator), generated from the unmodified reference like the others
CASES = ["cfg1_se1d", "se2d", "matern32_2d", "usa_f32", "se3d", "matern52_3d", "cfg2_full"]


def rel(a, b):
    a = torch.as_tensor(a).detach().cpu().to(torch.complex128)
    b = torch.as_tensor(b).detach().cpu().to(torch.complex128)
    return float(torch.linalg.norm((a - b).reshape(-1)) / (torch.linalg.norm(b.reshape(-1)) + 1e-300))


def build(case):
    import efgpnd
    from kernels import SquaredExponential, Matern
    g = np.load(os.path.join(GOLDEN, f"model_{case}.npz"))
    dt = torch.float32 if "float32" in str(g["dtype"]) else torch.float64
    x = torch.from_numpy(g["x"]).to(dt).cuda()
    y = torch.from_numpy(g["y"]).to(dt).cuda()
    d = x.shape[1]
    name = str(g["kernel"])
    if name == "se":
        k = SquaredExponential(dimension=d, init_lengthscale=float(g["ell"]), init_variance=float(g["var"]))
    else:
        k = Matern(dimension=d, init_lengthscale=float(g["ell"]), init_variance=float(g["var"]),
                   nu={"matern12": 0.5, "matern32": 1.5, "matern52": 2.5}[name])
    f64 = dt == torch.float64
    opts = {"cg_tolerance": 1e-10 if f64 else 1e-5, "max_cg_iterations": 2000}
    neps = 1e-12 if f64 else 1e-6
    m = efgpnd.EFGPND(x, y, k, sigmasq=float(g["sig2"]), eps=float(g["eps"]), nufft_eps=neps, opts=opts,
                      estimate_params=False)
    return efgpnd, m, g, dt, neps


@pytest.mark.parametrize("case", CASES)
def test_posterior_mean_and_variance(case):
    efgpnd, m, g, dt, neps = build(case)
    f64 = dt == torch.float64
    x_new = torch.from_numpy(g["x_new"]).to(dt).cuda()
    efgpnd.PROBE_DEVICE = "cpu"
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            mean, var = m.predict(x_new, return_variance=False)
            assert m.last_fit_stats["mtot"] == int(g["mtot"])
            assert abs(m.last_fit_stats["h"] - float(g["h"])) <= 1e-13 * float(g["h"])
            assert rel(m._ws, g["ws"]) < (1e-12 if f64 else 1e-6)
            # float32: seven runs against the reference's float32 golden (scripts/f32_errors.py, repeated suites) gave
            # beta 4e-5 ... 5.4e-4, mean ~1e-5, variances 0.5-4e-5, gradient 0.9-3.3e-4 -- beta moves by 10x from run to
            # run (the default spreading mode is not bit-reproducible and the float32 system is solved to 1e-5 only), so
            # the bars keep a wide margin over the worst value seen
            assert rel(m._beta, g["beta"]) < (1e-7 if f64 else 2e-3)
            assert rel(mean, g["mean"]) < (1e-6 if f64 else 1e-3)
            assert bool(torch.isnan(var).all())
            torch.manual_seed(321)
            _, var_s = m.predict(x_new, return_variance=True, variance_method="stochastic", hutchinson_probes=24)
            assert rel(var_s, g["var_stochastic"]) < (1e-5 if f64 else 2e-3)
            _, var_r = m.predict(x_new[:32], return_variance=True, variance_method="regular")
            assert rel(var_r, g["var_regular"]) < (1e-5 if f64 else 2e-3)
    finally:
        efgpnd.PROBE_DEVICE = None


@pytest.mark.parametrize("case", CASES)
def test_gradient_same_probe_seed(case):
    efgpnd, m, g, dt, neps = build(case)
    f64 = dt == torch.float64
    efgpnd.PROBE_DEVICE = "cpu"
    try:
        torch.manual_seed(1234)
        grads = m.compute_gradients(trace_samples=4, cg_tol=1e-10 if f64 else 1e-5, nufft_eps=neps,
                                    apply_gradients=True)
    finally:
        efgpnd.PROBE_DEVICE = None
    assert rel(grads, g["grad_T4"]) < (1e-5 if f64 else 5e-3)
    st = m.last_gradient_stats
    it = g["grad_T4_iters"]
    assert st["trace_num_rhs"] == int(it[2])
    if f64:
        assert abs(st["mean_cg_iters"] - int(it[0])) <= 2
        assert abs(st["trace_cg_iters"] - int(it[1])) <= 2
    else:        # float32 CG stops a few iterations earlier or later than the reference's float32 run (measured: <= 4 %)
        assert abs(st["mean_cg_iters"] - int(it[0])) <= max(2, int(it[0]) // 4)
        assert abs(st["trace_cg_iters"] - int(it[1])) <= max(2, int(it[1]) // 4)
    assert m._gp_params.raw.grad is not None and m._gp_params.raw.grad.shape == (3,)


def test_default_tolerances_match_reference_defaults():
    """Reference defaults (cg_tolerance 1e-4, nufft_eps 1e-4): mean within 10*eps-level agreement."""
    import efgpnd
    from kernels import SquaredExponential
    g = np.load(os.path.join(GOLDEN, "model_se2d.npz"))
    x = torch.from_numpy(g["x"]).cuda()
    y = torch.from_numpy(g["y"]).cuda()
    k = SquaredExponential(dimension=2, init_lengthscale=float(g["ell"]), init_variance=float(g["var"]))
    m = efgpnd.EFGPND(x, y, k, sigmasq=float(g["sig2"]), eps=float(g["eps"]), nufft_eps=1e-4, estimate_params=False)
    mean, _ = m.predict(torch.from_numpy(g["x_new"]).cuda(), return_variance=False)
    assert rel(mean, g["mean_default"]) < 2e-3


def test_refit_after_hyper_change_and_state_dict():
    efgpnd, m, g, dt, neps = build("se2d")
    x_new = torch.from_numpy(g["x_new"]).cuda()
    a, _ = m.predict(x_new, return_variance=False)
    assert "gp_params" in m.state_dict()
    m.kernel.set_hyper("lengthscale", 0.2)
    b, _ = m.predict(x_new, return_variance=False)
    assert rel(a, b) > 1e-3     # really refitted
    with pytest.raises(ValueError):
        m.predict(None)
    with pytest.raises(ValueError):
        m.predict(x_new, variance_method="bogus")


def test_cpu_tensors_fail_loudly():
    import efgpnd
    from efgp_b200._lib import EfgpError
    with pytest.raises(EfgpError):
        efgpnd.EFGPND(torch.zeros(10, 2), torch.zeros(10), "se", estimate_params=False)
