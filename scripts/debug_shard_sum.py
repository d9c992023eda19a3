"""Debug: is the sum of the shard transforms the transform of all points? (one GPU; explains 1-rank vs 2-rank beta)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gp-quadrature_b200"))
import torch
from efgp_b200 import ops

def rel(a, b): return float((a - b).norm() / b.norm())

for N in (300_000, 20_000_000):
    g = torch.Generator(device="cuda").manual_seed(17)
    x = torch.rand(N, 2, dtype=torch.float64, device="cuda", generator=g) * 2 - 1
    y = torch.sin(3 * x[:, 0]) * torch.cos(2 * x[:, 1]) + 0.1 * torch.randn(N, dtype=torch.float64, device="cuda", generator=g)
    h, m = 0.2, 35
    sg, sr = (4 * m + 1,) * 2, (2 * m + 1,) * 2
    for eps_rhs in (1e-4, 1e-9):
        full = ops.NufftPlan(x, h)
        vf, rf = full.type1_pair(y, sg, 6e-8, sr, eps_rhs)
        info_f = full.plan_info(sg, 6e-8) if hasattr(full, "plan_info") else None
        va = rb = None
        infos = []
        for lo, hi in ((0, N // 2), (N // 2, N)):
            p = ops.NufftPlan(x[lo:hi].contiguous(), h)
            v, r = p.type1_pair(y[lo:hi].contiguous(), sg, 6e-8, sr, eps_rhs)
            infos.append(p.plan_info(sg, 6e-8) if hasattr(p, "plan_info") else None)
            va = v if va is None else va + v
            rb = r if rb is None else rb + r
        # exact generator for reference at a few modes (direct sum over a subsample is not exact; use tight transform)
        tight = ops.NufftPlan(x, h).type1(None, sg, 1e-13)
        print(f"N={N} eps_rhs={eps_rhs}: gen shards-vs-full {rel(va, vf):.3e}  rhs shards-vs-full {rel(rb, rf):.3e}  "
              f"gen full-vs-tight {rel(vf, tight):.3e} shards-vs-tight {rel(va, tight):.3e}  info full {info_f} shards {infos}")
