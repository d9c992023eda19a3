"""Host-side cost of one bench step (cProfile over many small-N solves: the device work is negligible, what is
left is the Python / ctypes / launch path that bounds the step when the points are sharded 8 ways)."""
import os, sys, cProfile, pstats, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gp-quadrature_b200")); sys.path.insert(0, ROOT)
import torch, efgpnd, bench
from kernels.squared_exponential import SquaredExponential
N = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000
x, y = bench.synth(N, 2, 1, "cuda", torch.float64)
def solve():
    efgpnd._PLAN_CACHE.clear()
    kern = SquaredExponential(dimension=2, init_lengthscale=0.05, init_variance=1.0)
    m = efgpnd.EFGPND(x, y, kern, sigmasq=0.1, eps=1e-4, nufft_eps=1e-4, opts={"cg_tolerance": 1e-4}, estimate_params=False)
    return m.predict(x, return_variance=False)[0]
for _ in range(10): solve()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(50): solve()
torch.cuda.synchronize()
print(f"N={N}: {(time.perf_counter() - t0) / 50 * 1e3:.3f} ms per solve (wall, unprofiled)")
pr = cProfile.Profile()
pr.enable()
for _ in range(50): solve()
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(45)
st.sort_stats("cumulative").print_stats(45)
