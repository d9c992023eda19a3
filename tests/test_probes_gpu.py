"""Hutchinson probes streamed through the transforms (efgp_nufft_type1_probes / efgp_nufft_type2_probe_dots): the
Rademacher signs regenerated inside the spread kernel and inside the interpolation epilogue must be the SAME vectors
(exported explicitly by efgp_nufft_probes_export), so that

    type1_probes       == type1(Z)                       (efgpnd.py:183, F^* Z)
    type2_probe_dots   == sum_n Z * Re type2(modes)      (efgpnd.py:235-241)

to rounding, for every dimension / dtype / multi-probe kernel variant; and the gradient computed with the streamed
probes must equal the gradient computed from the exported probes through the stored-probe path."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = a.detach().double().reshape(-1).cpu(), b.detach().double().reshape(-1).cpu()
    return float(torch.linalg.norm(a - b) / (torch.linalg.norm(b) + 1e-300))


@pytest.mark.parametrize("d,dt,eps,n", [(1, torch.float64, 1e-6, 33), (2, torch.float64, 1e-4, 21), (2, torch.float64, 1e-6, 21),
                                        (2, torch.float64, 1e-12, 15), (3, torch.float64, 1e-5, 13), (2, torch.float32, 1e-4, 21),
                                        (3, torch.float32, 1e-3, 11)])
def test_streamed_probes_equal_explicit_probes(d, dt, eps, n):
    from efgp_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(10 + d)
    N = 60_000
    x = (torch.rand(N, d, dtype=torch.float64, device="cuda", generator=g) * 2 - 1).to(dt)
    plan = ops.NufftPlan(x, 0.37)
    carry = d != 1                                   # both counters: original index (float64 only) and slot
    plan.carry_index(carry)
    shape = (n,) * d
    seed, t0, nb = 123456789012345, 125, 7          # crosses a 128-probe block boundary; groups of 2/1 and 4/2/1
    z, sid = plan.probes_export(seed, t0, nb, shape, eps)
    if carry and dt == torch.float64:
        # counter = original index: the probes are the same for ANY sorted set of these points (other spacing, new sort)
        z_other, _ = ops.NufftPlan(x, 0.29).carry_index(True).probes_export(seed, t0, nb, shape, eps)
        assert torch.equal(z, z_other)
    assert z.shape == (nb, N) and bool(((z == 1) | (z == -1)).all())
    assert abs(float(z.double().mean())) < 0.02                       # unbiased signs
    assert float((z[0] * z[1]).double().mean().abs()) < 0.02          # different probes are uncorrelated
    z2, _ = plan.probes_export(seed, t0, nb, shape, eps, sort_id=sid)
    assert torch.equal(z, z2)                                         # a pure function of (seed, probe, slot)
    zo, _ = plan.probes_export(seed + 1, t0, nb, shape, eps, sort_id=sid)
    assert not torch.equal(z, zo)
    f64 = dt == torch.float64
    a, sid_a = plan.type1_probes(seed, t0, nb, shape, eps, sort_id=sid)
    b = plan.type1(z, shape, eps)
    assert sid_a == sid
    assert rel(a, b) < (1e-12 if f64 else 2e-5)
    scale = torch.rand(int(np.prod(shape)), dtype=dt, device="cuda", generator=g) + 0.5
    a_s, _ = plan.type1_probes(seed, t0, nb, shape, eps, scale=scale, sort_id=sid)
    assert rel(a_s, b * scale.reshape(shape)) < (1e-12 if f64 else 2e-5)
    modes = torch.randn((nb,) + shape, dtype=torch.complex128, device="cuda", generator=g).to(plan.cdtype)
    dots, _ = plan.type2_probe_dots(seed, t0, nb, modes, shape, eps, sort_id=sid)
    vals = plan.type2(modes, shape, eps, real_out=True)
    ref = (z.double() * vals.double()).sum(dim=1)
    scale_ref = float(vals.double().norm(dim=1).mean()) * np.sqrt(N)   # size of the sums' terms: |dots| itself may be small
    assert float((dots - ref).abs().max()) / scale_ref < (1e-12 if f64 else 1e-5)
    assert dots.dtype == torch.float64


def test_probe_transforms_refuse_a_changed_sort():
    """Slot counters (no original index in the records) tie the probes to ONE sorted set: a re-sort in between is an
    error, not a silently different estimator."""
    from efgp_b200 import ops
    from efgp_b200._lib import EfgpError
    g = torch.Generator(device="cuda").manual_seed(2)
    x = torch.rand(20_000, 2, dtype=torch.float64, device="cuda", generator=g)
    plan = ops.NufftPlan(x, 0.4)
    a, sid = plan.type1_probes(7, 0, 2, (15, 15), 1e-5)
    plan.ensure(0.41)                 # new spacing: the handle re-sorts
    modes = torch.randn(2, 15, 15, dtype=torch.complex128, device="cuda", generator=g)
    with pytest.raises(EfgpError):
        plan.type2_probe_dots(7, 0, 2, modes, (15, 15), 1e-5, sort_id=sid)


@pytest.mark.parametrize("d,kname", [(2, "se"), (3, "matern52"), (1, "se")])
def test_gradient_with_streamed_probes_equals_stored_probes(d, kname):
    """Same Philox probes, once regenerated inside the kernels (no (T, N) array) and once exported and pushed through
    the stored-probe path: the gradients agree to solver tolerance."""
    import efgpnd
    from kernels import SquaredExponential, Matern
    g = torch.Generator(device="cuda").manual_seed(31)
    N = 30_000
    x = torch.rand(N, d, dtype=torch.float64, device="cuda", generator=g) * 2 - 1
    y = torch.sin(3 * x[:, 0]) + 0.2 * torch.randn(N, dtype=torch.float64, device="cuda", generator=g)
    out = {}
    for mode in (True, "export"):
        k = SquaredExponential(dimension=d, init_lengthscale=0.4, init_variance=1.0) if kname == "se" else \
            Matern(dimension=d, init_lengthscale=0.8, init_variance=1.0, nu=2.5)
        # sigma^2 = 1: a well-conditioned system, so that CG converges long before its 2 M iteration cap and the
        # comparison is about the probes, not about where two rounding-different CG runs stopped
        m = efgpnd.EFGPND(x, y, k, sigmasq=1.0, eps=1e-3, opts={"device_probes": mode}, estimate_params=False)
        torch.manual_seed(77)
        out[mode] = m.compute_gradients(trace_samples=6, cg_tol=1e-9, nufft_eps=1e-10, apply_gradients=False).double().cpu()
        assert m.last_gradient_stats["trace_num_rhs"] == 12
        assert m.last_gradient_stats["trace_cg_iters"] < 2 * m.last_gradient_stats["feature_count"]
    assert float((out[True] - out["export"]).norm() / out["export"].norm()) < 1e-6
    # and a different seed gives a different Monte-Carlo estimate (the probes really enter)
    k = SquaredExponential(dimension=d, init_lengthscale=0.4, init_variance=1.0) if kname == "se" else \
        Matern(dimension=d, init_lengthscale=0.8, init_variance=1.0, nu=2.5)
    m = efgpnd.EFGPND(x, y, k, sigmasq=1.0, eps=1e-3, estimate_params=False)
    torch.manual_seed(78)
    other = m.compute_gradients(trace_samples=6, cg_tol=1e-9, nufft_eps=1e-10, apply_gradients=False).double().cpu()
    assert float((other - out[True]).norm()) > 0
