"""GPU timeline of the Toeplitz/CG phase of one bench step: every kernel between the last spread and the interp."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gp-quadrature_b200")); sys.path.insert(0, ROOT)
import torch, efgpnd, bench
from kernels.squared_exponential import SquaredExponential
from torch.profiler import profile, ProfilerActivity
N = int(float(sys.argv[1])) if len(sys.argv) > 1 else 12_500_000
x, y = bench.synth(N, 2, 1, "cuda", torch.float64)
def solve():
    efgpnd._PLAN_CACHE.clear()
    kern = SquaredExponential(dimension=2, init_lengthscale=0.05, init_variance=1.0)
    m = efgpnd.EFGPND(x, y, kern, sigmasq=0.1, eps=1e-4, nufft_eps=1e-4, opts={"cg_tolerance": 1e-4}, estimate_params=False)
    return m.predict(x, return_variance=False)[0]
for _ in range(4): solve()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    solve()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
t0 = evs[0].time_range.start
on = False
prev_end = None
for e in evs:
    s, d = e.time_range.start - t0, e.time_range.end - e.time_range.start
    if "interp2d" in e.name: break
    if on and d < 2000:
        print(f"{s:9.1f} {d:7.1f} gap {s - prev_end:6.1f}  {e.name[:90]}")
    if "spread2d_kernel<double, 5" in e.name: on = True
    prev_end = s + d
