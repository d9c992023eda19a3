import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "gp-quadrature_b200"))
import torch
from efgp_b200 import ops
N = int(float(sys.argv[1])) if len(sys.argv) > 1 else 10_000_000
m = int(sys.argv[2]) if len(sys.argv) > 2 else 21
torch.manual_seed(0)
x = torch.rand(N, 3, dtype=torch.float64, device="cuda") * 2 - 1
y = torch.randn(N, dtype=torch.float64, device="cuda")
plan = ops.NufftPlan(x, 0.4117)
mt = 2 * m + 1
for _ in range(2):
    v = plan.type1(None, (4 * m + 1,) * 3, 6e-8)
    Fy = plan.type1(y, (mt,) * 3, 1e-4)
torch.cuda.synchronize()
print("ok", float(Fy.abs().mean()))
