"""One logdet_slq call on the config-2 grid (N = 1e5, 2-D Matern-3/2, 100 probes x 25 steps) for an ncu launch list:

    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/lanczos_launches.csv \
        python scripts/lanczos_profile.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gp-quadrature_b200"))
import efgpnd  # noqa: E402
from kernels import Matern  # noqa: E402

torch.manual_seed(0)
N = 100_000
x = torch.rand(N, 2, dtype=torch.float64, device="cuda") * 2 - 1
y = torch.sin(3 * x[:, 0]) + 0.2 * torch.randn(N, dtype=torch.float64, device="cuda")
m = efgpnd.EFGPND(x, y, Matern(dimension=2, init_lengthscale=0.2, init_variance=1.0, nu=1.5), sigmasq=0.04, eps=1e-4,
                  nufft_eps=1e-4, estimate_params=False)
m._compute_common_parameters()
torch.cuda.synchronize()
torch.cuda.nvtx.range_push("logdet_slq")
print(efgpnd.logdet_slq(m._ws, 0.04, m._toeplitz, probes=100, steps=25, dtype=torch.float64, n=N))
torch.cuda.nvtx.range_pop()
