"""TEST INFRASTRUCTURE ONLY -- exact NUDFT stand-in for ``pytorch_finufft.functional``.

The reference (`/root/reference/efgpnd.py:11`) imports the third-party, un-vendored
`pytorch-finufft` / `finufft` wheels (unpinned in `/root/reference/setup.py:17-18`), which are
not installable in this image.  FINUFFT's contract is "the nonuniform DFT below, to relative
l2 error ~eps"; this module evaluates that DFT *exactly* (dense, O(N*M), chunked), so anything
within 10*eps of it is within 10*eps (+ one eps) of FINUFFT.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / reference arm may
import this.  The product (`gp-quadrature_b200/`) never does.

Semantics restated from the reference call sites:
  * `efgpnd.py:1496-1499`  finufft_type1(points(d,N), values(...,N), output_shape, eps=, isign=-1, modeord=False)
  * `efgpnd.py:1533-1536, 1546-1549` finufft_type2(points, targets(...,*modes), eps=, isign=+1, modeord=False)
  * `efgpnd.py:1679` finufft_type2(..., modeord=True)  (FFT ordering of the mode axes)
  * definition of F: `utils/gradient_tests.py:37-39`  F = exp(2*pi*i x xi^T), phi = 2*pi*h*x (`efgpnd.py:1451`)

    type1: out[..., k] = sum_n values[..., n] * exp(isign * i * <k, points[:, n]>)
    type2: out[..., n] = sum_k targets[..., k] * exp(isign * i * <k, points[:, n]>)
  per-axis integer modes k = -(n//2) .. (n-1)//2 (centred), or FFT order if modeord is truthy.
"""
import torch

_CHUNK_ELEMS = 1 << 24  # cap on the dense (chunk x modes) phase matrix, in complex elements


def _cdtype(rdtype):
    return torch.complex64 if rdtype == torch.float32 else torch.complex128


def _mode_vector(n, modeord, dtype):
    k = torch.arange(-(n // 2), (n - 1) // 2 + 1, dtype=dtype)
    if modeord:
        k = torch.fft.ifftshift(k)
    return k


def _axis_phases(points, shape, isign, modeord):
    """List of per-axis phase matrices E_a[n, k_a] = exp(isign*i*k_a*points[a, n]) in complex128."""
    out = []
    for a, n in enumerate(shape):
        k = _mode_vector(int(n), modeord, torch.float64)
        ang = points[a].to(torch.float64)[:, None] * k[None, :]
        out.append(torch.polar(torch.ones_like(ang), float(isign) * ang))
    return out


def _dense_rows(E):
    """Kronecker-combine per-axis phases into (n_chunk, prod(shape)) with C-order modes."""
    F = E[0]
    for Ea in E[1:]:
        F = (F[:, :, None] * Ea[:, None, :]).reshape(F.shape[0], -1)
    return F


def finufft_type1(points, values, output_shape, eps=1e-6, isign=-1, modeord=False, **_):
    if points.ndim == 1:
        points = points[None, :]
    d, N = points.shape
    if isinstance(output_shape, int):
        output_shape = (output_shape,)
    output_shape = tuple(int(s) for s in output_shape)
    assert len(output_shape) == d
    cd = _cdtype(points.dtype)
    batch = values.shape[:-1]
    v = values.reshape(-1, N).to(torch.complex128)
    M = 1
    for s in output_shape:
        M *= s
    out = None     # accumulated functionally (no in-place ops) so the stand-in also runs under torch.vmap,
    step = max(1, _CHUNK_ELEMS // max(M, 1))   # which pg_classifier.py:357-358 wraps around type1 / type2
    for s0 in range(0, N, step):
        sl = slice(s0, min(N, s0 + step))
        F = _dense_rows(_axis_phases(points[:, sl], output_shape, isign, modeord))
        part = v[:, sl] @ F
        out = part if out is None else out + part
    if out is None:
        out = torch.zeros(v.shape[0], M, dtype=torch.complex128)
    return out.reshape(*batch, *output_shape).to(cd)


def finufft_type2(points, targets, eps=1e-6, isign=+1, modeord=False, **_):
    if points.ndim == 1:
        points = points[None, :]
    d, N = points.shape
    shape = tuple(targets.shape[-d:])
    batch = targets.shape[:-d]
    cd = _cdtype(points.dtype)
    M = 1
    for s in shape:
        M *= s
    f = targets.reshape(-1, M).to(torch.complex128)
    parts = []
    step = max(1, _CHUNK_ELEMS // max(M, 1))
    for s0 in range(0, N, step):
        sl = slice(s0, min(N, s0 + step))
        F = _dense_rows(_axis_phases(points[:, sl], shape, isign, modeord))
        parts.append(f @ F.T)
    out = torch.cat(parts, dim=1) if parts else torch.empty(f.shape[0], 0, dtype=torch.complex128)
    return out.reshape(*batch, N).to(cd)
