"""TEST INFRASTRUCTURE ONLY -- fast CPU NUFFT with the pytorch_finufft.functional call signature
(finufft_type1 / finufft_type2), standing in for FINUFFT where the exact O(N M) NUDFT of
oracle/pytorch_finufft is too slow: the timed CPU baseline in bench.py and large-N parity checks.

Algorithm (restated from the FINUFFT paper, the library itself is absent here -- see cnufft.c):
spread with the exponential-of-semicircle kernel on a sigma=2 fine grid (C + OpenMP, cnufft.c),
FFT (numpy/pocketfft), deconvolve by the kernel's Fourier transform and crop; type 2 is the adjoint.
Call sites it serves: efgpnd.py:1496-1499 (type 1, isign=-1), :1533-1549 (type 2, isign=+1),
:1679 (type 2, modeord=True).  Only tests/, __graft_entry__.smoke() and bench.py may import it.
"""
import ctypes
import math
import os

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def _lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "libcnufft.so")
        if not os.path.exists(path):
            raise RuntimeError(f"{path} missing: run `make -C oracle`")
        L = ctypes.CDLL(path)
        dp = ctypes.POINTER(ctypes.c_double)
        ip = ctypes.POINTER(ctypes.c_int64)
        L.cn_spread.restype = None
        L.cn_spread.argtypes = [ctypes.c_int, ctypes.c_int64, dp, dp, ctypes.c_int, ctypes.c_int, dp, ip, dp,
                                ctypes.c_int]
        L.cn_interp.restype = None
        L.cn_interp.argtypes = [ctypes.c_int, ctypes.c_int64, dp, ctypes.c_int, ctypes.c_int, dp, ip, dp, dp,
                                ctypes.c_int]
        L.cn_num_threads.restype = ctypes.c_int
        _LIB = L
    return _LIB


def num_threads():
    return int(_lib().cn_num_threads())


NTHREADS = 0  # 0 = all OpenMP threads


def _dp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def _next_smooth_even(n):
    n = max(int(n), 2)
    while True:
        m = n
        for p in (2, 3, 5):
            while m % p == 0:
                m //= p
        if m == 1 and n % 2 == 0:
            return n
        n += 1


_KER_CACHE = {}


def _kernel(eps):
    """width, beta and the Horner table: coef[i, j] multiplies s^j for footprint slot i."""
    eps = max(float(eps), 1e-15)
    key = eps
    if key in _KER_CACHE:
        return _KER_CACHE[key]
    w = int(min(16, max(2, math.ceil(math.log10(1.0 / eps)) + 1)))
    beta = 2.30 * w
    deg = w + 4
    # Chebyshev nodes in s; slot i covers distances d = t + i - w/2, t = (s+1)/2
    q = np.cos(np.pi * (np.arange(deg + 1) + 0.5) / (deg + 1))
    coef = np.zeros((w, deg + 1))
    for i in range(w):
        z = ((q + 1) / 2 + i - w / 2) / (w / 2)
        f = np.exp(beta * (np.sqrt(np.clip(1 - z * z, 0, None)) - 1))
        c = np.polynomial.chebyshev.chebfit(q, f, deg)
        coef[i] = np.polynomial.chebyshev.cheb2poly(c)
    _KER_CACHE[key] = (w, beta, np.ascontiguousarray(coef))
    return _KER_CACHE[key]


def _psihat(w, beta, nf, kmax):
    """Fourier transform of psi(x) = phi(2x/w) at frequencies k/nf, k = 0..kmax (Gauss-Legendre)."""
    nodes, wts = np.polynomial.legendre.leggauss(4 * w + 30)
    xq = nodes * (w / 2)
    f = np.exp(beta * (np.sqrt(1 - nodes ** 2) - 1)) * wts * (w / 2)
    k = np.arange(kmax + 1)[:, None]
    return (f[None, :] * np.cos(2 * np.pi * k * xq[None, :] / nf)).sum(1)


def _mode_index(n, modeord):
    k = np.arange(-(n // 2), (n - 1) // 2 + 1)
    if modeord:
        k = np.fft.ifftshift(k)
    return k


def _prep(points, shape, eps):
    pts = points.detach().cpu().to(torch.float64).numpy()
    if pts.ndim == 1:
        pts = pts[None, :]
    d, N = pts.shape
    w, beta, coef = _kernel(eps)
    nf = np.array([_next_smooth_even(max(2 * n, 2 * w)) for n in shape], dtype=np.int64)
    X = np.ascontiguousarray(pts / (2 * np.pi) * nf[:, None].astype(np.float64))
    return d, N, w, beta, coef, nf, X


def finufft_type1(points, values, output_shape, eps=1e-6, isign=-1, modeord=False, **_):
    """out[..., k] = sum_n values[..., n] exp(isign i k . points[:, n])"""
    shape = tuple(int(s) for s in output_shape)
    d, N, w, beta, coef, nf, X = _prep(points, shape, eps)
    vals = values.detach().cpu()
    cd = torch.complex64 if points.dtype == torch.float32 else torch.complex128
    lead = tuple(vals.shape[:-1])
    v = np.ascontiguousarray(vals.to(torch.complex128).reshape(-1, N).numpy())
    L = _lib()
    out = np.empty((v.shape[0],) + shape, dtype=np.complex128)
    dec = [1.0 / _psihat(w, beta, int(nf[a]), shape[a] // 2) for a in range(d)]
    idx = [_mode_index(shape[a], modeord) for a in range(d)]
    for b in range(v.shape[0]):
        grid = np.zeros(tuple(int(n) for n in nf), dtype=np.complex128)
        L.cn_spread(d, N, _dp(X), _dp(v[b].view(np.float64)), w, coef.shape[1], _dp(coef),
                    nf.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)), _dp(grid.view(np.float64)), NTHREADS)
        # sum_l g[l] e^{isign 2 pi i k l / nf}: forward FFT for isign < 0, unnormalised inverse otherwise
        spec = np.fft.fftn(grid) if isign < 0 else np.fft.ifftn(grid) * grid.size
        sub = spec
        for a in range(d):
            sub = np.take(sub, idx[a] % nf[a], axis=a)
            sh = [1] * d
            sh[a] = -1
            sub = sub * dec[a][np.abs(idx[a])].reshape(sh)
        out[b] = sub
    return torch.from_numpy(out).reshape(lead + shape).to(cd)


def finufft_type2(points, targets, eps=1e-6, isign=+1, modeord=False, **_):
    """out[..., n] = sum_k targets[..., k] exp(isign i k . points[:, n])"""
    pts_d = points.shape[0] if points.ndim > 1 else 1
    shape = tuple(int(s) for s in targets.shape[-pts_d:])
    d, N, w, beta, coef, nf, X = _prep(points, shape, eps)
    cd = torch.complex64 if points.dtype == torch.float32 else torch.complex128
    lead = tuple(targets.shape[:-d])
    f = targets.detach().cpu().to(torch.complex128).reshape((-1,) + shape).numpy()
    L = _lib()
    out = np.empty((f.shape[0], N), dtype=np.complex128)
    dec = [1.0 / _psihat(w, beta, int(nf[a]), shape[a] // 2) for a in range(d)]
    idx = [_mode_index(shape[a], modeord) for a in range(d)]
    for b in range(f.shape[0]):
        sub = f[b]
        for a in range(d):
            sh = [1] * d
            sh[a] = -1
            sub = sub * dec[a][np.abs(idx[a])].reshape(sh)
        spec = np.zeros(tuple(int(n) for n in nf), dtype=np.complex128)
        spec[np.ix_(*[idx[a] % nf[a] for a in range(d)])] = sub
        # g[l] = sum_k spec[k] e^{isign 2 pi i k l / nf}
        grid = np.ascontiguousarray(np.fft.ifftn(spec) * spec.size if isign > 0 else np.fft.fftn(spec))
        L.cn_interp(d, N, _dp(X), w, coef.shape[1], _dp(coef), nf.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)),
                    _dp(grid.view(np.float64)), _dp(out[b].view(np.float64)), NTHREADS)
    return torch.from_numpy(out).reshape(lead + (N,)).to(cd)
