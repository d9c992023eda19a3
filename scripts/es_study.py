"""Empirical accuracy of the ES-kernel NUFFT for candidate (W, sigma, beta factor, polynomial degree): relative l2
error of a 2-D type-1 against the exact NUDFT (numpy, float64, CPU).  Mirrors csrc/es_kernel.cpp: beta =
f * pi * W * (1 - 1/(2 sigma)), piecewise polynomial of degree `deg` per grid offset (Chebyshev fit), deconvolution by
Gauss-Legendre quadrature of the kernel's Fourier transform.  Used to choose the width policy."""
import sys
import numpy as np


def phi(beta, z):
    return np.exp(beta * (np.sqrt(np.clip(1 - z * z, 0, None)) - 1)) * (np.abs(z) <= 1)


def build(W, beta, deg):
    # coefficient table for offsets i = 0..W-1 in s = 2u, u in [-1/2, 1/2)
    q = np.cos(np.pi * (np.arange(deg + 1) + 0.5) / (deg + 1))
    coef = np.zeros((W, deg + 1))
    for i in range(W):
        z = (i - 0.5 * (W - 1) - 0.5 * q) / (0.5 * W)
        c = np.polynomial.chebyshev.chebfit(q, phi(beta, z), deg)
        coef[i] = np.polynomial.chebyshev.cheb2poly(c)
    return coef


def psihat(W, beta, nf, k):
    nodes, wts = np.polynomial.legendre.leggauss(128)
    f = phi(beta, nodes) * wts
    return 0.5 * W * (f[None, :] * np.cos(np.pi * W * (k[:, None] / nf) * nodes[None, :])).sum(1)


def weights(coef, W, X):
    """X fine-grid coordinate -> (first index, W weights) with the library's odd/even cell convention."""
    odd = W & 1
    Xc = X + (0.5 if odd else 0.0)
    c = np.floor(Xc).astype(int)
    s = 2 * (Xc - c) - 1            # in [-1, 1)
    OFF = W // 2 - (0 if odd else 1)
    P = np.stack([np.polyval(coef[i][::-1], s) for i in range(W)], 1)
    return c - OFF, P


def type1_2d(x, c, n, sigma, W, bf, deg, exact_kernel=False):
    nf = int(np.ceil(sigma * n / 2) * 2)
    beta = bf * np.pi * W * (1 - 1 / (2 * sigma))
    coef = build(W, beta, deg)
    grid = np.zeros((nf, nf))
    X = (x % 1.0) * nf
    i0, P0 = weights(coef, W, X[:, 0])
    i1, P1 = weights(coef, W, X[:, 1])
    for a in range(W):
        for b in range(W):
            np.add.at(grid, ((i0 + a) % nf, (i1 + b) % nf), c * P0[:, a] * P1[:, b])
    spec = np.fft.fft2(grid)
    k = np.arange(-(n // 2), (n - 1) // 2 + 1)
    ph = psihat(W, beta, nf, np.abs(k).astype(float))
    # (grid point l = c - OFF + i sits at distance i - (W-1)/2 - u from X for either parity: no phase correction)
    return spec[np.ix_(k % nf, k % nf)] / (ph[:, None] * ph[None, :])


def exact_2d(x, c, n):
    k = np.arange(-(n // 2), (n - 1) // 2 + 1)
    E0 = np.exp(-2j * np.pi * np.outer(k, x[:, 0]))
    E1 = np.exp(-2j * np.pi * np.outer(k, x[:, 1]))
    return (E0 * c[None, :]) @ E1.T


if __name__ == "__main__":
    rng = np.random.default_rng(0)
    N = 3000
    x = rng.random((N, 2))
    cases = {"gen": (141, 6e-8, np.ones(N)), "rhs": (71, 1e-4, rng.standard_normal(N))}
    which = sys.argv[1] if len(sys.argv) > 1 else "gen"
    n, eps, c = cases[which]
    ref = exact_2d(x, c, n)
    print(f"case {which}: n={n} eps={eps}")
    for sigma in (2.0, 2.5, 3.0, 4.0, 6.0, 8.0):
        for W in range(3, 11):
            for bf in (0.93, 0.97, 1.0):
                for deg in (W - 1, W, W + 1, W + 2):
                    out = type1_2d(x, c, n, sigma, W, bf, deg)
                    err = np.linalg.norm(out - ref) / np.linalg.norm(ref)
                    if err < 30 * eps and err > eps / 300:
                        print(f"sigma {sigma:3.1f} W {W:2d} bf {bf:4.2f} deg {deg:2d} err {err:9.2e} ({err / eps:6.2f} eps)")
