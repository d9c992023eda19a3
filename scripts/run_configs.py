"""Run BASELINE.json's other configurations at full size once (timing + sanity; parity is covered by tests/).
Prints one JSON line per configuration."""
import os, sys, time, json, io, contextlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gp-quadrature_b200")); sys.path.insert(0, ROOT)
import torch, efgpnd
from kernels import SquaredExponential, Matern

def sync(): torch.cuda.synchronize()
def timed(fn, reps=1):
    sync(); t0 = time.perf_counter()
    for _ in range(reps): out = fn()
    sync(); return (time.perf_counter() - t0) / reps * 1e3, out

def field(x):
    return torch.sin(3 * x[:, 0]) * torch.cos(4 * x[:, 1]) + (0.3 * torch.sin(5 * x[:, 2]) if x.shape[1] > 2 else 0)

which = sys.argv[1:] or ["2", "4", "5"]
dev = "cuda"
g = torch.Generator(device=dev).manual_seed(0)
for c in which:
    efgpnd._PLAN_CACHE.clear()
    torch.cuda.empty_cache()
    if c == "2":   # 2-D Matern-3/2 posterior mean + variance, N = 1e5
        N, d = 100_000, 2
        x = torch.rand(N, d, dtype=torch.float64, device=dev, generator=g) * 2 - 1
        y = field(x) + 0.2 * torch.randn(N, dtype=torch.float64, device=dev, generator=g)
        k = Matern(dimension=d, init_lengthscale=0.2, init_variance=1.0, nu=1.5)
        m = efgpnd.EFGPND(x, y, k, sigmasq=0.04, eps=1e-4, nufft_eps=1e-4, estimate_params=False)
        with contextlib.redirect_stdout(io.StringIO()):
            m.predict(x[:1000], return_variance=True, hutchinson_probes=100)     # warm-up
            t_fit, _ = timed(lambda: m.predict(x, return_variance=False, force_recompute=True))
            t_var, (mean, var) = timed(lambda: m.predict(x, return_variance=True, hutchinson_probes=1000))
        print(json.dumps({"config": 2, "N": N, "d": d, "kernel": "matern32", "mtot": m.last_fit_stats["mtot"],
                          "fit+mean_ms": t_fit, "mean+stochastic_var_1000probes_ms": t_var,
                          "mean_cg_iters": m.last_fit_stats["mean_cg_iters"], "var_min": float(var.min()),
                          "var_mean": float(var.mean())}), flush=True)
    if c == "4":   # 3-D SE gradient step, N = 1e7, 64 probes
        N, d, T = 10_000_000, 3, 64
        x = torch.rand(N, d, dtype=torch.float64, device=dev, generator=g) * 2 - 1
        y = field(x) + 0.2 * torch.randn(N, dtype=torch.float64, device=dev, generator=g)
        k = SquaredExponential(dimension=d, init_lengthscale=0.1, init_variance=1.0)
        m = efgpnd.EFGPND(x, y, k, sigmasq=0.04, eps=1e-4, estimate_params=False)
        torch.manual_seed(1234)
        t1, g1 = timed(lambda: m.compute_gradients(trace_samples=T, cg_tol=1e-5, nufft_eps=1e-5, apply_gradients=False))
        torch.manual_seed(1234)
        t2, g2 = timed(lambda: m.compute_gradients(trace_samples=T, cg_tol=1e-5, nufft_eps=1e-5, apply_gradients=False))
        st = m.last_gradient_stats
        print(json.dumps({"config": 4, "N": N, "d": d, "T": T, "mtot": st["mtot"], "first_step_ms": t1, "step_ms": t2,
                          "mean_cg_iters": st["mean_cg_iters"], "trace_cg_iters": st["trace_cg_iters"],
                          "grad": [float(v) for v in g2], "mem_GB": torch.cuda.max_memory_allocated() / 1e9}), flush=True)
    if c == "5":   # 3-D Matern-5/2, N = 2e7, eps 1e-6, fp64: hyper-learning steps (3 Adam iterations here)
        N, d = 20_000_000, 3
        x = torch.rand(N, d, dtype=torch.float64, device=dev, generator=g) * 2 - 1
        y = field(x) + 0.2 * torch.randn(N, dtype=torch.float64, device=dev, generator=g)
        k = Matern(dimension=d, init_lengthscale=0.5, init_variance=1.0, nu=2.5)
        m = efgpnd.EFGPND(x, y, k, sigmasq=0.04, eps=1e-6, estimate_params=False)
        with contextlib.redirect_stdout(io.StringIO()):
            t, _ = timed(lambda: m.optimize_hyperparameters(max_iters=3, lr=0.1, trace_samples=10))
        st = m.last_gradient_stats
        print(json.dumps({"config": 5, "N": N, "d": d, "adam_iters": 3, "T": 10, "mtot_last": st["mtot"], "total_ms": t,
                          "ms_per_iter_incl_final_refit": t / 3, "trace_cg_iters": st["trace_cg_iters"],
                          "lengthscale": float(m.kernel.get_hyper("lengthscale")),
                          "mem_GB": torch.cuda.max_memory_allocated() / 1e9}), flush=True)
    del x, y, m
