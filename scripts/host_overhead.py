"""Where does the fixed (N-independent) part of a solve go?  Runs the bench's solve at small N with
host timers around the phases."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gp-quadrature_b200")); sys.path.insert(0, ROOT)
import torch, efgpnd
from kernels.squared_exponential import SquaredExponential
import bench
N = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100000
x, y = bench.synth(N, 2, 1, "cuda", torch.float64)
opts = {"cg_tolerance": 1e-4}
def solve():
    efgpnd._PLAN_CACHE.clear()
    t0 = time.perf_counter()
    kern = SquaredExponential(dimension=2, init_lengthscale=0.05, init_variance=1.0)
    t1 = time.perf_counter()
    m = efgpnd.EFGPND(x, y, kern, sigmasq=0.1, eps=1e-4, nufft_eps=1e-4, opts=opts, estimate_params=False)
    t2 = time.perf_counter()
    m._compute_common_parameters()
    torch.cuda.synchronize(); t3 = time.perf_counter()
    mean, _ = m.predict(x, return_variance=False)
    torch.cuda.synchronize(); t4 = time.perf_counter()
    return [t1 - t0, t2 - t1, t3 - t2, t4 - t3]
for _ in range(5): solve()
import numpy as np
ts = np.array([solve() for _ in range(20)]) * 1e3
print("ms: kernel ctor %.3f | EFGPND ctor %.3f | fit %.3f | predict %.3f | total %.3f" % (*ts.mean(0), ts.mean(0).sum()))
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for _ in range(20): solve()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(35)
