"""GPU timeline of one bench step (torch.profiler / CUPTI): kernels in start order with idle gaps."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gp-quadrature_b200")); sys.path.insert(0, ROOT)
import torch, efgpnd, bench
from kernels.squared_exponential import SquaredExponential
from torch.profiler import profile, ProfilerActivity
N = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000_000
x, y = bench.synth_range(0, N, 2, torch.device("cuda"), torch.float64)
def solve():
    efgpnd._PLAN_CACHE.clear()
    efgpnd._BBOX_CACHE.clear()
    kern = SquaredExponential(dimension=2, init_lengthscale=0.05, init_variance=1.0)
    m = efgpnd.EFGPND(x, y, kern, sigmasq=0.1, eps=1e-4, nufft_eps=1e-4, opts={"cg_tolerance": 1e-4}, estimate_params=False)
    return m.predict(x, return_variance=False)[0]
for _ in range(4): solve()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    solve(); solve()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
t0 = evs[0].time_range.start
prev_end = t0
rows = []
for e in evs:
    gap = e.time_range.start - prev_end
    rows.append((e.time_range.start - t0, e.time_range.end - e.time_range.start, gap, e.name[:70]))
    prev_end = max(prev_end, e.time_range.end)
half = len(rows) // 2
TH = float(sys.argv[2]) if len(sys.argv) > 2 else 15.0
print("second solve: start_us dur_us gap_before_us name   (only kernels >20us or gaps > TH us)")
base = rows[half][0]
tot_gap = 0
for s, d, g, n in rows[half:]:
    if g > 0: tot_gap += g
    if d > 20 or g > TH:
        print(f"{s-base:10.1f} {d:9.1f} {g:8.1f}  {n}")
print("kernels in second solve:", len(rows) - half, " total idle gaps us:", tot_gap, " span us:", rows[-1][0] + rows[-1][1] - base)
