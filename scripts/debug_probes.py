"""Debug: streamed vs exported probes inside the gradient (which partial sums differ?)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gp-quadrature_b200")); sys.path.insert(0, ROOT)
import torch
import efgpnd
from efgp_b200 import ops
from kernels import SquaredExponential

g = torch.Generator(device="cuda").manual_seed(31)
N, d = 30_000, 2
x = torch.rand(N, d, dtype=torch.float64, device="cuda", generator=g) * 2 - 1
y = torch.sin(3 * x[:, 0]) + 0.2 * torch.randn(N, dtype=torch.float64, device="cuda", generator=g)
for rep in range(2):
    for mode in (True, "export", False):
        k = SquaredExponential(dimension=d, init_lengthscale=0.4, init_variance=1.0)
        m = efgpnd.EFGPND(x, y, k, sigmasq=0.1, eps=1e-3, opts={"device_probes": mode}, estimate_params=False)
        torch.manual_seed(77)
        gr = m.compute_gradients(trace_samples=6, cg_tol=1e-10, nufft_eps=1e-10, apply_gradients=False)
        st = m.last_gradient_stats
        print(rep, mode, [f"{float(v):.10f}" for v in gr], st["mean_cg_iters"], st["trace_cg_iters"], st["mtot"])
# operator level at the same settings
plan = ops.NufftPlan(x, 0.3)
for eps, n in ((1e-10, 11), (1e-10, 9), (1e-3, 11)):
    shape = (n, n)
    z, sid = plan.probes_export(5, 0, 6, shape, eps)
    a, _ = plan.type1_probes(5, 0, 6, shape, eps, sort_id=sid)
    b = plan.type1(z, shape, eps)
    modes = torch.randn((6,) + shape, dtype=torch.complex128, device="cuda", generator=g)
    dots, _ = plan.type2_probe_dots(5, 0, 6, modes, shape, eps, sort_id=sid)
    vals = plan.type2(modes, shape, eps, real_out=True)
    ref = (z * vals).sum(1)
    print(eps, n, float((a - b).abs().max() / b.abs().max()), (dots - ref).abs().max().item(), ref.abs().max().item())
