"""A/B of the real-FFT (Hermitian) CG matvec against the general one on the batched 3-D / 2-D systems of configs 2, 4:

    python scripts/hermitian_cg_bench.py >> gpurun_out/hermitian_cg.jsonl

Rows are ws F^* Z for real +-1 point probes Z (the kernel-derivative rows of the gradient, efgpnd.py:183-197)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gp-quadrature_b200"))
import efgpnd  # noqa: E402
from efgp_b200 import ops  # noqa: E402
from kernels import SquaredExponential, Matern  # noqa: E402


def run(tag, N, d, kernel, eps, rows, s2, tol):
    torch.manual_seed(0)
    x = torch.rand(N, d, dtype=torch.float64, device="cuda") * 2 - 1
    y = torch.sin(3 * x[:, 0]) + 0.2 * torch.randn(N, dtype=torch.float64, device="cuda")
    m = efgpnd.EFGPND(x, y, kernel, sigmasq=s2, eps=eps, nufft_eps=eps, estimate_params=False)
    m._compute_common_parameters()
    mt = m.last_fit_stats["mtot"]
    ws = m._ws.real.contiguous()
    Z = torch.randint(0, 2, (rows, N), device="cuda").to(torch.float64) * 2 - 1
    op = efgpnd.NUFFT(x, torch.zeros(d, dtype=torch.float64, device="cuda"), m.last_fit_stats["h"], eps)
    b = op.type1(Z, (mt,) * d, scale=ws).reshape(rows, -1)
    del Z
    plan = m._toeplitz.plan
    minv = 1.0 / (N * ws * ws + s2)
    out = {"case": tag, "N": N, "d": d, "M": int(ws.numel()), "fft_shape": list(plan.fft_shape), "rows": rows}
    sols = {}
    for name, herm in (("general", False), ("hermitian", True), ("general2", False), ("hermitian2", True)):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        xs, it = plan.cg(b, None, ops.MODE_A_MEAN, ws, s2, tol, minv=minv, x0_zero=True, hermitian=herm)
        e1.record()
        torch.cuda.synchronize()
        out[name + "_ms"] = e0.elapsed_time(e1)
        out[name + "_iters"] = it
        sols[name] = xs
    out["rel_diff"] = float(torch.linalg.norm(sols["hermitian"] - sols["general"]) / torch.linalg.norm(sols["general"]))
    out["speedup"] = out["general2_ms"] / out["hermitian2_ms"]
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    which = sys.argv[1:] or ["2", "4"]
    if "2" in which:
        run("config2_grid", 100_000, 2, Matern(dimension=2, init_lengthscale=0.2, init_variance=1.0, nu=1.5), 1e-4, 256, 0.04, 1e-5)
    if "4" in which:
        run("config4_grid", 10_000_000, 3, SquaredExponential(dimension=3, init_lengthscale=0.1, init_variance=1.0), 1e-4, 64, 0.04, 1e-5)
