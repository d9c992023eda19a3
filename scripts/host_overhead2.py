import os, sys, time, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gp-quadrature_b200")); sys.path.insert(0, ROOT)
import torch, efgpnd, cg as cgmod
from kernels.squared_exponential import SquaredExponential
import bench
N = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100000
x, y = bench.synth(N, 2, 1, "cuda", torch.float64)
acc = collections.defaultdict(float)
def wrap(obj, name, label, sync=False):
    f = getattr(obj, name)
    def g(*a, **k):
        t = time.perf_counter(); r = f(*a, **k)
        if sync: torch.cuda.synchronize()
        acc[label] += time.perf_counter() - t; return r
    setattr(obj, name, g)
SYNC = len(sys.argv) > 2
wrap(efgpnd, "get_xis", "get_xis", SYNC)
wrap(efgpnd, "_domain_length", "domain_length(aminmax+item)", SYNC)
wrap(efgpnd, "compute_convolution_vector_vectorized_dD", "conv_vector(sort+spread9)", SYNC)
wrap(efgpnd.NUFFT, "type1", "type1(y)", SYNC)
wrap(efgpnd.NUFFT, "type2", "type2", SYNC)
wrap(efgpnd.ToeplitzND, "__init__", "toeplitz_init", SYNC)
wrap(cgmod.ConjugateGradients, "solve", "cg_solve", SYNC)
wrap(SquaredExponential, "spectral_density", "spectral_density", SYNC)
wrap(efgpnd, "create_jacobi_precond", "jacobi", SYNC)
def solve():
    efgpnd._PLAN_CACHE.clear()
    kern = SquaredExponential(dimension=2, init_lengthscale=0.05, init_variance=1.0)
    m = efgpnd.EFGPND(x, y, kern, sigmasq=0.1, eps=1e-4, nufft_eps=1e-4, opts={"cg_tolerance": 1e-4}, estimate_params=False)
    mean, _ = m.predict(x, return_variance=False)
    return mean
for _ in range(5): solve()
torch.cuda.synchronize(); acc.clear()
t0 = time.perf_counter()
for _ in range(20): solve()
torch.cuda.synchronize()
tot = (time.perf_counter() - t0) / 20 * 1e3
print(f"N={N} sync={SYNC} total {tot:.3f} ms/solve")
for k, v in acc.items(): print(f"   {k:32s} {v/20*1e3:.3f} ms")
print(f"   other {tot - sum(acc.values())/20*1e3:.3f} ms")
