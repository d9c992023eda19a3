"""Ad-hoc timing of the NUFFT stages at large N (not the bench)."""
import sys, os, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "gp-quadrature_b200"))
import torch
from efgp_b200 import ops

def timeit(fn, n=3):
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); r = fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts), r

N = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 2
dt = torch.float64 if (len(sys.argv) <= 3 or sys.argv[3] == "f64") else torch.float32
m = int(sys.argv[4]) if len(sys.argv) > 4 else 35
h = float(sys.argv[5]) if len(sys.argv) > 5 else 0.4515
torch.manual_seed(0)
x = torch.rand(N, d, dtype=dt, device="cuda") * 2 - 1
y = torch.randn(N, dtype=dt, device="cuda")
mtot = 2 * m + 1
t0 = time.time()
plan = ops.NufftPlan(x, h)
t, v = timeit(lambda: plan.type1(None, (4 * m + 1,) * d, 6e-8), 1)
print(f"first conv type1 (incl sort) {t:.2f} ms")
t, v = timeit(lambda: plan.type1(None, (4 * m + 1,) * d, 6e-8))
print(f"conv type1 (ones, eps 6e-8) {t:.3f} ms  -> {N/t/1e3:.0f} Mpts/s")
t, Fy = timeit(lambda: plan.type1(y, (mtot,) * d, 1e-4))
print(f"type1 y (eps 1e-4, shared grid) {t:.3f} ms -> {N/t/1e3:.0f} Mpts/s")
t, mean = timeit(lambda: plan.type2(Fy, (mtot,) * d, 1e-4, real_out=True))
print(f"type2 real (eps 1e-4, shared grid) {t:.3f} ms -> {N/t/1e3:.0f} Mpts/s")
plan2 = ops.NufftPlan(x, h)
t, Fy2 = timeit(lambda: plan2.type1(y, (mtot,) * d, 1e-4), 1)
print(f"standalone type1 y first (incl sort) {t:.2f} ms")
t, Fy2 = timeit(lambda: plan2.type1(y, (mtot,) * d, 1e-4))
print(f"standalone type1 y (own grid) {t:.3f} ms -> {N/t/1e3:.0f} Mpts/s")
t, _ = timeit(lambda: plan2.type2(Fy2, (mtot,) * d, 1e-4, real_out=True))
print(f"standalone type2 (own grid) {t:.3f} ms -> {N/t/1e3:.0f} Mpts/s")
print("rel diff Fy shared vs own", float((Fy - Fy2).norm() / Fy2.norm()))
