"""CPU-only checks of the product's host-side logic (no CUDA): the eps-driven frequency grid of
gp-quadrature_b200/utils/kernels.py against the reference's golden rows (tests/golden/host.npz, reference
utils/kernels.py:72-143), the L-independent split of its bisections, and the point-shard arithmetic."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

NAMES = ["se", "matern12", "matern32", "matern52"]


def _kernel(name, d, ell, var):
    from kernels import SquaredExponential, Matern
    if name == "se":
        return SquaredExponential(dimension=d, init_lengthscale=ell, init_variance=var)
    nu = {"matern12": 0.5, "matern32": 1.5, "matern52": 2.5}[name]
    return Matern(dimension=d, init_lengthscale=ell, init_variance=var, nu=nu)


def test_product_get_xis_matches_reference_rows():
    from utils.kernels import get_xis
    rows = np.load(os.path.join(GOLDEN, "host.npz"))["get_xis"]
    for row in rows[::2]:
        ki, d, ell, eps, L, h, mtot, x0, x1 = row
        k = _kernel(NAMES[int(ki)], int(d), float(ell), 1.3)
        xis, hh, mt = get_xis(k, float(eps), float(L), use_integral=True)
        assert mt == int(mtot)
        assert abs(hh - h) <= 1e-12 * abs(h)
        assert abs(float(xis[0]) - x0) <= 1e-10 * abs(x0)


@pytest.mark.parametrize("name,d", [("se", 1), ("se", 2), ("matern32", 2), ("matern52", 3)])
def test_truncation_bounds_split_is_equivalent(name, d):
    """EFGPND runs the two bisections before it knows the domain length (they do not depend on it)."""
    from utils.kernels import get_xis, truncation_bounds
    k = _kernel(name, d, 0.21, 0.8)
    b = truncation_bounds(k, 1e-5)
    for L in (0.3, 2.0, 17.5):
        x_a, h_a, m_a = get_xis(k, 1e-5, L, use_integral=True)
        x_b, h_b, m_b = get_xis(k, 1e-5, L, use_integral=True, bounds=b)
        assert (h_a, m_a) == (h_b, m_b)
        assert torch.equal(x_a, x_b)


def test_shard_bounds_partition_the_points():
    import efgpnd
    for n, w in ((10, 3), (100_000_000, 8), (7, 8), (0, 2)):
        cuts = [efgpnd.shard_bounds(n, w, r) for r in range(w)]
        assert cuts[0][0] == 0 and cuts[-1][1] == n
        assert all(cuts[r][1] == cuts[r + 1][0] for r in range(w - 1))
        sizes = [hi - lo for lo, hi in cuts]
        assert max(sizes) - min(sizes) <= 1
