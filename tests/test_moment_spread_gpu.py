"""The dense-point-set spreading paths of csrc/spread.cu against the exact nonuniform DFT (oracle, tests only):

  * spread2d_moment_kernel + moment_to_grid_kernel: per-cell MOMENTS instead of footprints (fp64, 2-D, widths 7..9,
    >= 48 points per fine-grid cell), all-ones / real / complex strengths, batched;
  * spread2d_moment_pair_kernel: the generator F^*1 and the right-hand side F^*y of a fit in ONE pass over the records
    (efgp_nufft_type1_pair), one batched R2C.

Each test also asserts, through the library's stage report, that the path under test is the one that ran (the small
point sets of tests/test_ops_gpu.py are sparse on their grids and take the register-footprint kernels)."""
import numpy as np
import pytest
import torch

from conftest import oracle

pytestmark = pytest.mark.gpu

O = oracle()
from pytorch_finufft import functional as nudft  # noqa: E402  (oracle stand-in, tests only)


def rel(a, b):
    a = torch.as_tensor(a).detach().cpu().reshape(-1).to(torch.complex128)
    b = torch.as_tensor(b).detach().cpu().reshape(-1).to(torch.complex128)
    return float(torch.linalg.norm(a - b) / torch.linalg.norm(b))


def points(N, seed, clustered=False):
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(N, 2, dtype=torch.float64, generator=g) * 2 - 1
    if clustered:      # half of the points in a small blob: long runs in a few cells, short runs elsewhere
        x[: N // 2] = 0.05 * torch.randn(N // 2, 2, dtype=torch.float64, generator=g) + 0.3
    return x


def stages_of(fn):
    from efgp_b200 import _lib
    _lib.timing_enable(True)
    out = fn()
    torch.cuda.synchronize()
    st = _lib.timing_report()
    _lib.timing_enable(False)
    return out, st


# For small mode counts the grid-choice cost model takes a large upsampling factor (the grid is nearly free), so the
# widths below are those of sigma = 4 .. 8: eps 6e-8 -> 7, 1e-9 -> 9 (es_error_model); N is large enough for >= 48
# points per cell on a (8 n)^2 grid.
@pytest.mark.parametrize("n,eps,N", [(9, 6e-8, 300_000), (9, 1e-9, 300_000), (12, 1e-9, 500_000), (15, 3e-8, 700_000)])
@pytest.mark.parametrize("kind", ["ones", "real", "complex"])
def test_moment_spread_matches_exact_nudft(n, eps, N, kind):
    from efgp_b200 import ops
    h = 0.41
    x = points(N, 3, clustered=(n == 12))
    g = torch.Generator().manual_seed(5)
    if kind == "ones":
        c, cref = None, torch.ones(N, dtype=torch.complex128)
    elif kind == "real":
        c = torch.randn(N, dtype=torch.float64, generator=g)
        cref = c.to(torch.complex128)
    else:
        c = torch.randn(N, dtype=torch.float64, generator=g) + 1j * torch.randn(N, dtype=torch.float64, generator=g)
        cref = c
    plan = ops.NufftPlan(x.cuda(), h)
    out, st = stages_of(lambda: plan.type1(None if c is None else c.cuda(), (n, n), eps))
    assert any(k == "type1.moment_to_grid" for k in st), sorted(st)          # the moment path ran
    phi = (2 * np.pi * h * x).T.contiguous()
    ref = nudft.finufft_type1(phi, cref, (n, n), isign=-1, modeord=False)
    assert rel(out, ref) < 10 * eps


def test_moment_spread_batched_rows_are_independent():
    from efgp_b200 import ops
    N, n, eps, h = 300_000, 9, 6e-8, 0.37
    x = points(N, 7)
    g = torch.Generator().manual_seed(8)
    C = torch.randn(3, N, dtype=torch.float64, generator=g)
    plan = ops.NufftPlan(x.cuda(), h)
    out, st = stages_of(lambda: plan.type1(C.cuda(), (n, n), eps))
    assert "type1.moment_to_grid" in st
    for b in range(3):
        one = plan.type1(C[b].cuda(), (n, n), eps)
        assert rel(out[b], one) < 1e-13


@pytest.mark.parametrize("m,eps_gen,eps_rhs,N", [(2, 1e-9, 1e-4, 400_000), (3, 1e-9, 1e-4, 800_000), (3, 6e-8, 1e-4, 800_000)])
def test_fused_generator_and_rhs_pass_matches_exact_nudft(m, eps_gen, eps_rhs, N):
    """efgp_nufft_type1_pair on a dense set: generator on (4 m + 1)^2 at eps_gen (width 9 or 7) and ws * F^* y on
    (2 m + 1)^2 at eps_rhs (width 5 on the generator's grid) from one pass; both against the exact sums."""
    from efgp_b200 import ops
    h = 0.33
    x = points(N, 11)
    g = torch.Generator().manual_seed(12)
    y = torch.randn(N, dtype=torch.float64, generator=g)
    sg, sr = (4 * m + 1,) * 2, (2 * m + 1,) * 2
    scale = (0.5 + torch.rand(sr[0] * sr[1], dtype=torch.float64, generator=g))
    plan = ops.NufftPlan(x.cuda(), h)
    (v, r), st = stages_of(lambda: plan.type1_pair(y.cuda(), sg, eps_gen, sr, eps_rhs, scale=scale.cuda()))
    assert any(k.startswith("spread2d_pair_w") for k in st), sorted(st)        # the fused pass ran
    phi = (2 * np.pi * h * x).T.contiguous()
    vref = nudft.finufft_type1(phi, torch.ones(N, dtype=torch.complex128), sg, isign=-1, modeord=False)
    rref = nudft.finufft_type1(phi, y.to(torch.complex128), sr, isign=-1, modeord=False).reshape(-1) * scale
    assert rel(v, vref) < 10 * eps_gen
    assert rel(r, rref) < 10 * eps_rhs
    # and against the two separate transforms of the same plan (register / moment kernels, own passes)
    v2 = plan.type1(None, sg, eps_gen)
    r2 = plan.type1(y.cuda(), sr, eps_rhs).reshape(-1) * scale.cuda()
    assert rel(v, v2) < 1e-12
    assert rel(r, r2) < 1e-12
