"""Device timings of the SURVEY.md section 8(f) rows (run on the GPU box):

    python scripts/next_rows_bench.py > gpurun_out/next_rows.jsonl

  logdet_slq            library Lanczos (efgp_lanczos) vs the same recurrence written as torch ops around the
                        batched matvec (what the module did before), reference defaults probes=100, steps=25
  pg sigma_apply        B right-hand sides through type-1 -> batched CG on the weighted Toeplitz plan -> type-2
  sample_gp_spectral    one batched type-2 at N points
"""
import json
import math
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gp-quadrature_b200"))
sys.path.insert(0, os.path.join(ROOT, "gp-quadrature_b200", "polyagamma_classification"))

import efgpnd  # noqa: E402
from efgp_b200 import ops  # noqa: E402
from kernels import SquaredExponential, Matern  # noqa: E402
import pg_operators as pg  # noqa: E402
import vanilla_gp_sampling as vs  # noqa: E402


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best, out


def torch_loop_slq(ws, s2, toeplitz, probes, steps, dtype, n):
    """The recurrence as torch ops around the batched fused matvec (one host sync per step)."""
    dev = toeplitz.device
    wr = ws.real.to(dtype)
    cd = efgpnd._cmplx(dtype)
    A = efgpnd._FusedOperator(wr.to(cd), toeplitz, s2, ops.MODE_A_VAR, cd)
    z = efgpnd._rademacher_like_reference((probes, wr.numel()), dtype, dev)
    znorm = z.norm(dim=1)
    q = (z / znorm[:, None]).to(cd)
    q_prev = torch.zeros_like(q)
    beta_prev = torch.zeros(probes, dtype=dtype, device=dev)
    al, be = [], []
    alive = torch.ones(probes, dtype=torch.bool, device=dev)
    for _ in range(steps):
        v = A(q) - beta_prev[:, None] * q_prev
        alpha = torch.sum(q.conj() * v, dim=1).real
        v = v - alpha[:, None] * q
        beta = torch.linalg.norm(v, dim=1).real
        al.append(torch.where(alive, alpha, torch.ones_like(alpha)))
        be.append(torch.where(alive, beta, torch.zeros_like(beta)))
        alive = alive & (beta >= 1e-12)
        q_prev, beta_prev = q, torch.where(alive, beta, torch.zeros_like(beta))
        q = torch.where(alive[:, None], v / beta.clamp_min(1e-300)[:, None], q)
        if not bool(alive.any()):
            break
    a, b = torch.stack(al, 1), torch.stack(be, 1)
    T = torch.diag_embed(a) + torch.diag_embed(b[:, :-1], 1) + torch.diag_embed(b[:, :-1], -1)
    ev, U = torch.linalg.eigh(T)
    quad = (U[:, 0, :] ** 2 * torch.log(ev.clamp_min(1e-18))).sum(1) * znorm ** 2
    return float(quad.mean()) + n * math.log(s2)


def bench_logdet(tag, N, d, kernel, eps, sig2):
    torch.manual_seed(0)
    x = torch.rand(N, d, dtype=torch.float64, device="cuda") * 2 - 1
    y = torch.sin(3 * x[:, 0]) + 0.2 * torch.randn(N, dtype=torch.float64, device="cuda")
    m = efgpnd.EFGPND(x, y, kernel, sigmasq=sig2, eps=eps, nufft_eps=eps, estimate_params=False)
    m._compute_common_parameters()
    kw = dict(probes=100, steps=25, dtype=torch.float64, n=N)
    torch.manual_seed(1)
    t_lib, v_lib = timed(lambda: efgpnd.logdet_slq(m._ws, sig2, m._toeplitz, **kw))
    torch.manual_seed(1)
    t_loop, v_loop = timed(lambda: torch_loop_slq(m._ws, sig2, m._toeplitz, 100, 25, torch.float64, N))
    print(json.dumps({"row": "logdet_slq", "case": tag, "N": N, "d": d, "M": int(m._ws.numel()),
                      "fft_shape": list(m._toeplitz.fft_shape), "probes": 100, "steps": 25,
                      "library_lanczos_ms": t_lib, "torch_op_loop_ms": t_loop, "value": v_lib}), flush=True)


def bench_pg(N, B):
    torch.manual_seed(0)
    X = torch.rand(N, 2, dtype=torch.float64, device="cuda")
    kern = SquaredExponential(dimension=2, init_lengthscale=0.1, init_variance=1.0)
    t0 = time.perf_counter()
    sp = pg._build_spectral_state(X, kern, spectral_eps=1e-4, trunc_eps=1e-4, nufft_eps=1e-6, rdtype=torch.float64,
                                  cdtype=torch.complex128, device=torch.device("cuda"))
    torch.cuda.synchronize()
    t_state = (time.perf_counter() - t0) * 1e3
    delta = 0.05 + 0.2 * torch.rand(N, dtype=torch.float64, device="cuda")
    t_wt, wt = timed(lambda: pg._build_weighted_toeplitz(delta, sp))
    sig, info = pg._make_sigma_apply(sp, delta, cg_tol=1e-6, weighted_toeplitz=wt)
    z = torch.randn(B, N, dtype=torch.float64, device="cuda")
    t_sig, _ = timed(lambda: sig(z))
    print(json.dumps({"row": "pg_operators", "N": N, "d": 2, "mtot": sp.mtot, "rhs": B, "spectral_state_ms": t_state,
                      "weighted_toeplitz_ms": t_wt, "sigma_apply_ms": t_sig, "cg_iters": info["cg_iters"]}), flush=True)


def bench_sampler(N, ns):
    torch.manual_seed(0)
    x = torch.rand(N, 2, dtype=torch.float64, device="cuda") * 2 - 1
    t, s = timed(lambda: vs.sample_gp_spectral_approx(x, num_samples=ns, length_scale=0.05, variance=1.0,
                                                      spectral_eps=1e-4, trunc_eps=1e-4, nufft_eps=1e-6, seed=3))
    print(json.dumps({"row": "sample_gp_spectral_approx", "N": N, "d": 2, "samples": ns, "ms": t,
                      "Mpts_per_s": N * ns / t / 1e3, "sample_std": float(s.std())}), flush=True)


if __name__ == "__main__":
    bench_logdet("config2_grid", 100_000, 2, Matern(dimension=2, init_lengthscale=0.2, init_variance=1.0, nu=1.5), 1e-4, 0.04)
    bench_logdet("config4_grid", 10_000_000, 3, SquaredExponential(dimension=3, init_lengthscale=0.1, init_variance=1.0), 1e-4, 0.04)
    bench_pg(1_000_000, 8)
    bench_sampler(10_000_000, 4)
