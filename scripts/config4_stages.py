import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gp-quadrature_b200")); sys.path.insert(0, ROOT)
import torch, efgpnd
from efgp_b200 import _lib
from kernels import SquaredExponential
N, d, T = 10_000_000, 3, 64
g = torch.Generator(device="cuda").manual_seed(0)
x = torch.rand(N, d, dtype=torch.float64, device="cuda", generator=g) * 2 - 1
y = torch.sin(3 * x[:, 0]) * torch.cos(4 * x[:, 1]) + 0.3 * torch.sin(5 * x[:, 2]) + 0.2 * torch.randn(N, dtype=torch.float64, device="cuda", generator=g)
k = SquaredExponential(dimension=d, init_lengthscale=0.1, init_variance=1.0)
m = efgpnd.EFGPND(x, y, k, sigmasq=0.04, eps=1e-4, estimate_params=False)
torch.manual_seed(1234)
m.compute_gradients(trace_samples=T, cg_tol=1e-5, nufft_eps=1e-5, apply_gradients=False)
torch.cuda.synchronize()
_lib.timing_enable(True)
torch.manual_seed(1234)
t0 = time.perf_counter()
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
    m.compute_gradients(trace_samples=T, cg_tol=1e-5, nufft_eps=1e-5, apply_gradients=False)
    torch.cuda.synchronize()
print("step ms", (time.perf_counter() - t0) * 1e3)
st = _lib.timing_report()
tot = 0
for k_, v in st.items():
    print(f"  {k_:34s} {v['ms']:9.2f} ms  scopes {v['scopes']}")
    tot += v["ms"]
print("  library stages total", tot)
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=14, max_name_column_width=60))
