"""Torch-facing wrappers over the C ABI: device tensors in, device tensors out.

PyTorch is used only for device memory and streams; all arithmetic on the O(N) and O(M log M)
paths happens in libefgp_b200.so.
"""
import contextlib
import ctypes

import torch

from . import _lib
from ._lib import check, i64_array, dbl_array

MODE_T, MODE_GV, MODE_A_MEAN, MODE_A_VAR = 0, 1, 2, 3


def _dtype_code(real_dtype):
    if real_dtype == torch.float32:
        return 0
    if real_dtype == torch.float64:
        return 1
    raise TypeError(f"unsupported dtype {real_dtype}; EFGP runs in float32 or float64")


def cmplx_of(real_dtype):
    return torch.complex64 if real_dtype == torch.float32 else torch.complex128


def real_of(dtype):
    if dtype in (torch.complex64, torch.float32):
        return torch.float32
    return torch.float64


def _require_cuda(t, what):
    if not t.is_cuda:
        raise _lib.EfgpError(f"{what} must be a CUDA tensor: the EFGP hot path has no CPU fallback")


_NULL = contextlib.nullcontext()


def _on(dev):
    """Device guard for the C calls; entering torch.cuda.device costs ~10 us of host time per call, which is on the
    critical path of the small launch-bound sections, so it is skipped when `dev` is already current."""
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    return _NULL if torch.cuda.current_device() == idx else torch.cuda.device(idx)


def _stream(dev):
    return ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


class NufftPlan:
    """Bin-sorted point set + fine-grid plans for x (N, d) scaled by h (phi = 2 pi h (x - xcen))."""

    def __init__(self, x, h, xcen=None):
        _require_cuda(x, "x")
        if x.ndim == 1:
            x = x[:, None]
        self.x = x.contiguous()
        self.N, self.d = self.x.shape
        self.rdtype = self.x.dtype
        self.cdtype = cmplx_of(self.rdtype)
        self.device = self.x.device
        self.h = float(h)
        xc = [0.0] * self.d if xcen is None else [float(v) for v in xcen]
        self._h = ctypes.c_void_p()
        check(_lib.lib().efgp_nufft_create(ctypes.byref(self._h), int(self.d), _dtype_code(self.rdtype)))
        with _on(self.device):
            check(_lib.lib().efgp_nufft_set_points(self._h, _ptr(self.x), int(self.N), self.h, dbl_array(xc),
                                                   _stream(self.device)))

        self.xcen = tuple(xc)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _lib.lib().efgp_nufft_destroy(h)
            except Exception:
                pass

    def carry_index(self, on=True):
        """Sorted sets carry the points' original indices (efgp_nufft_carry_index): the regenerated probes then do not
        depend on the sort.  Changing the setting drops the existing sorted sets."""
        check(_lib.lib().efgp_nufft_carry_index(self._h, 1 if on else 0))
        return self

    def ensure(self, h, xcen=None):
        """Same points, (possibly) new spacing / centre (efgp_nufft_rescale): a no-op when nothing changed."""
        h = float(h)
        xc = self.xcen if xcen is None else tuple(float(v) for v in xcen)
        if h != self.h or xc != self.xcen:
            check(_lib.lib().efgp_nufft_rescale(self._h, h, dbl_array(xc)))
            self.h, self.xcen = h, xc
        return self

    def _scale(self, scale, M):
        if scale is None:
            return None
        if scale.is_complex():
            scale = scale.real
        scale = scale.to(device=self.device, dtype=self.rdtype).reshape(-1).contiguous()
        if scale.numel() != M:
            raise ValueError(f"scale has {scale.numel()} entries, expected {M}")
        return scale

    def type1(self, vals, out_shape, eps, isign=-1, modeord=False, scale=None):
        """vals (N,) or (B, N), real or complex, or None for all-ones -> (…, *out_shape) complex."""
        out_shape = tuple(int(s) for s in out_shape)
        if len(out_shape) != self.d:
            raise ValueError(f"out_shape {out_shape} does not match dimension {self.d}")
        M = 1
        for s in out_shape:
            M *= s
        batched = vals is not None and vals.ndim > 1
        if vals is None:
            B, is_c, v = 1, 0, None
        else:
            _require_cuda(vals, "vals")
            if vals.shape[-1] != self.N:
                raise ValueError(f"vals has trailing size {vals.shape[-1]}, expected {self.N}")
            is_c = 1 if vals.is_complex() else 0
            B = 1
            for sdim in vals.shape[:-1]:
                B *= int(sdim)
            v = vals.to(self.cdtype if is_c else self.rdtype).reshape(B, self.N).contiguous()
        sc = self._scale(scale, M)
        out = torch.empty((B,) + out_shape, dtype=self.cdtype, device=self.device)
        with _on(self.device):
            check(_lib.lib().efgp_nufft_type1(self._h, _ptr(v), is_c, B, i64_array(out_shape), float(eps),
                                              int(isign), int(bool(modeord)), _ptr(sc), _ptr(out),
                                              _stream(self.device)))
        if batched:
            return out.reshape(tuple(vals.shape[:-1]) + out_shape)
        return out.reshape(out_shape)

    def type1_pair(self, y, shape_gen, eps_gen, shape_rhs, eps_rhs, scale=None):
        """(F^*1 on shape_gen at eps_gen, scale * F^*y on shape_rhs at eps_rhs) with one sort that carries y."""
        shape_gen = tuple(int(s) for s in shape_gen)
        shape_rhs = tuple(int(s) for s in shape_rhs)
        if len(shape_gen) != self.d or len(shape_rhs) != self.d:
            raise ValueError("shapes do not match the dimension")
        _require_cuda(y, "y")
        if y.is_complex() or y.numel() != self.N:
            raise ValueError("y must be a real (N,) tensor")
        yv = y.to(self.rdtype).reshape(self.N).contiguous()
        M = 1
        for s in shape_rhs:
            M *= s
        sc = self._scale(scale, M)
        out_gen = torch.empty(shape_gen, dtype=self.cdtype, device=self.device)
        out_rhs = torch.empty(shape_rhs, dtype=self.cdtype, device=self.device)
        with _on(self.device):
            check(_lib.lib().efgp_nufft_type1_pair(self._h, _ptr(yv), i64_array(shape_gen), float(eps_gen),
                                                   i64_array(shape_rhs), float(eps_rhs), _ptr(sc), _ptr(out_gen),
                                                   _ptr(out_rhs), _stream(self.device)))
        return out_gen, out_rhs

    def type2(self, fk, shape, eps, isign=+1, modeord=False, scale=None, real_out=False):
        """fk (*shape) / (M,) or batched (B, …) complex -> (N,) or (B, N), real if real_out."""
        shape = tuple(int(s) for s in shape)
        if len(shape) != self.d:
            raise ValueError(f"shape {shape} does not match dimension {self.d}")
        M = 1
        for s in shape:
            M *= s
        _require_cuda(fk, "fk")
        f = fk.to(self.cdtype)
        if f.numel() % M != 0:
            raise ValueError(f"mode array of {f.numel()} elements is not a multiple of {M}")
        B = f.numel() // M
        batched = not (tuple(f.shape) == shape or tuple(f.shape) == (M,))
        f = f.reshape(B, M).contiguous()
        sc = self._scale(scale, M)
        out = torch.empty((B, self.N), dtype=self.rdtype if real_out else self.cdtype, device=self.device)
        with _on(self.device):
            check(_lib.lib().efgp_nufft_type2(self._h, _ptr(f), B, i64_array(shape), float(eps), int(isign),
                                              int(bool(modeord)), _ptr(sc), 1 if real_out else 0, _ptr(out),
                                              _stream(self.device)))
        return out if batched else out.reshape(self.N)


def _probe_methods():
    """Hutchinson probes streamed through the transforms (efgp_nufft_type1_probes / _type2_probe_dots / _probes_export)."""

    def type1_probes(self, seed, t0, nb, out_shape, eps, scale=None, sort_id=0):
        """(scale * F^* Z_t for t0 <= t < t0 + nb, sort_id): (nb, *out_shape) complex; Z regenerated in the kernel."""
        out_shape = tuple(int(s) for s in out_shape)
        M = 1
        for s_ in out_shape:
            M *= s_
        sc = self._scale(scale, M)
        out = torch.empty((int(nb),) + out_shape, dtype=self.cdtype, device=self.device)
        sid = ctypes.c_uint64(int(sort_id))
        with _on(self.device):
            check(_lib.lib().efgp_nufft_type1_probes(self._h, ctypes.c_uint64(int(seed)), int(t0), int(nb),
                                                     i64_array(out_shape), float(eps), _ptr(sc), _ptr(out),
                                                     ctypes.byref(sid), _stream(self.device)))
        return out, int(sid.value)

    def type2_probe_dots(self, seed, t0, nb, modes, shape, eps, scale=None, sort_id=0):
        """sums[b] = sum_n Z_{t0+b}[n] Re(F (scale * modes[b]))[n]: (nb,) float64; nothing of size N is written."""
        shape = tuple(int(s) for s in shape)
        M = 1
        for s_ in shape:
            M *= s_
        f = modes.to(self.cdtype).reshape(int(nb), M).contiguous()
        sc = self._scale(scale, M)
        sums = torch.empty(int(nb), dtype=torch.float64, device=self.device)
        sid = ctypes.c_uint64(int(sort_id))
        with _on(self.device):
            check(_lib.lib().efgp_nufft_type2_probe_dots(self._h, ctypes.c_uint64(int(seed)), int(t0), int(nb), _ptr(f),
                                                         i64_array(shape), float(eps), _ptr(sc), _ptr(sums),
                                                         ctypes.byref(sid), _stream(self.device)))
        return sums, int(sid.value)

    def probes_export(self, seed, t0, nb, shape, eps, sort_id=0):
        """The probes themselves, (nb, N) real in original point order (tests, A/B runs)."""
        shape = tuple(int(s) for s in shape)
        z = torch.empty((int(nb), self.N), dtype=self.rdtype, device=self.device)
        sid = ctypes.c_uint64(int(sort_id))
        with _on(self.device):
            check(_lib.lib().efgp_nufft_probes_export(self._h, ctypes.c_uint64(int(seed)), int(t0), int(nb),
                                                      i64_array(shape), float(eps), _ptr(z), ctypes.byref(sid),
                                                      _stream(self.device)))
        return z, int(sid.value)

    return type1_probes, type2_probe_dots, probes_export


NufftPlan.type1_probes, NufftPlan.type2_probe_dots, NufftPlan.probes_export = _probe_methods()


class LazyIters:
    """Iteration count of a solve that was only enqueued (efgp_cg_solve_async): int(obj) waits for the stream to
    deliver the two convergence counters into pinned host memory and applies the reference's counting rule."""
    _ring, _owners, _next = None, [None] * 64, 0

    def __init__(self, device, max_iter, batched, early):
        cls = LazyIters
        if cls._ring is None:
            cls._ring = torch.zeros((64, 2), dtype=torch.int32).pin_memory()
        slot = cls._next
        cls._next = (slot + 1) % 64
        prev = cls._owners[slot]
        if prev is not None:
            prev.resolve()                       # the slot is about to be reused
        cls._owners[slot] = self
        self._buf = cls._ring[slot]
        self._slot = slot
        self._device, self._max_iter, self._batched, self._early = device, int(max_iter), bool(batched), bool(early)
        self._event = None
        self._value = None

    @property
    def ptr(self):
        return ctypes.c_void_p(self._buf.data_ptr())

    def enqueued(self):
        self._event = torch.cuda.Event()
        self._event.record(torch.cuda.current_stream(self._device))

    def set(self, value):
        self._value = int(value)
        LazyIters._owners[self._slot] = None

    def resolve(self):
        if self._value is None:
            if self._event is None:          # the call that should have enqueued the solve raised: nothing will arrive
                self.set(self._max_iter)
                return self._value
            self._event.synchronize()
            active, jmax = int(self._buf[0]), int(self._buf[1])
            if self._early and active <= 0:
                v = min(jmax + 2, self._max_iter) if self._batched else jmax + 1
            else:
                v = self._max_iter
            self.set(v)
        return self._value

    def __int__(self):
        return self.resolve()

    __index__ = __int__

    def __eq__(self, other):
        return int(self) == other

    def __hash__(self):
        return hash(int(self))

    def __repr__(self):
        return str(int(self))


class ToeplitzPlan:
    """Cached spectrum of the zero-padded F^H F generator v ((4m+1)^d complex)."""

    def __init__(self, v, force_pow2=True):
        _require_cuda(v, "v")
        if not v.is_complex():
            v = v.to(cmplx_of(v.dtype))
        self.v = v.contiguous()
        self.cdtype = self.v.dtype
        self.rdtype = real_of(self.cdtype)
        self.device = v.device
        self.Ls = list(self.v.shape)
        self.d = len(self.Ls)
        self.ns = [(L + 1) // 2 for L in self.Ls]
        self.size = 1
        for n in self.ns:
            self.size *= n
        self._h = ctypes.c_void_p()
        with _on(self.device):
            check(_lib.lib().efgp_toeplitz_create(ctypes.byref(self._h), _ptr(self.v), self.d, i64_array(self.Ls),
                                                  _dtype_code(self.rdtype), 1 if force_pow2 else 0,
                                                  _stream(self.device)))
        shp = (ctypes.c_int64 * self.d)()
        check(_lib.lib().efgp_toeplitz_fft_shape(self._h, shp))
        self.fft_shape = [int(s) for s in shp]

    @classmethod
    def from_handle(cls, handle, v):
        """Wrap a Toeplitz handle the library created itself (efgp_fit); v is the generator it was built from."""
        self = cls.__new__(cls)
        self.v = v
        self.cdtype = v.dtype
        self.rdtype = real_of(v.dtype)
        self.device = v.device
        self.Ls = list(v.shape)
        self.d = len(self.Ls)
        self.ns = [(L + 1) // 2 for L in self.Ls]
        self.size = 1
        for n in self.ns:
            self.size *= n
        self._h = handle
        shp = (ctypes.c_int64 * self.d)()
        check(_lib.lib().efgp_toeplitz_fft_shape(self._h, shp))
        self.fft_shape = [int(s) for s in shp]
        return self

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _lib.lib().efgp_toeplitz_destroy(h)
            except Exception:
                pass

    def _ws(self, ws):
        if ws is None:
            return None
        if ws.is_complex():
            ws = ws.real
        return ws.to(device=self.device, dtype=self.rdtype).reshape(-1).contiguous()

    def apply(self, x, mode=MODE_T, ws=None, sigmasq=0.0):
        """x (…, M) complex -> same shape."""
        _require_cuda(x, "x")
        xs = x.to(self.cdtype).reshape(-1, self.size).contiguous()
        y = torch.empty_like(xs)
        w = self._ws(ws)
        with _on(self.device):
            check(_lib.lib().efgp_toeplitz_apply(self._h, _ptr(xs), xs.shape[0], _ptr(w), float(sigmasq), int(mode),
                                                 _ptr(y), _stream(self.device)))
        return y.reshape(x.shape)

    def cg(self, b, x0, mode, ws, sigmasq, tol, max_iter=None, early_stopping=True, minv=None, batched=None,
           x0_zero=False, hermitian=False, lazy_iters=False):
        """Solve A x = b for the fused operator `mode`; returns (x, iters_completed).  x0_zero=True promises that
        x0 is all zeros (x0 may then be None): the library skips the matvec A x0.  hermitian=True promises
        b[-k] = conj(b[k]) (and the same for x0): the matvec then runs on real FFTs of half the size.
        lazy_iters=True: the second result is a LazyIters; when the solve runs as one cooperative kernel the call
        does not synchronise and the count is read only when somebody asks for it."""
        _require_cuda(b, "b")
        if batched is None:
            batched = b.ndim > 1
        bs = b.to(self.cdtype).reshape(-1, self.size).contiguous()
        if x0_zero:
            x = torch.zeros_like(bs)
        else:
            x = x0.to(self.cdtype).reshape(-1, self.size).clone().contiguous()
        w = self._ws(ws)
        mi = self._ws(minv)
        iters = ctypes.c_int(0)
        if lazy_iters:
            mi_host = 2 * self.size if max_iter is None else int(max_iter)
            lz = LazyIters(self.device, mi_host, batched, early_stopping)
            with _on(self.device):
                check(_lib.lib().efgp_cg_solve_async(self._h, _ptr(w), float(sigmasq), int(mode), _ptr(mi), _ptr(bs),
                                                     _ptr(x), bs.shape[0], float(tol), mi_host,
                                                     1 if early_stopping else 0, 1 if batched else 0,
                                                     1 if x0_zero else 0, 1 if hermitian else 0, lz.ptr,
                                                     ctypes.byref(iters), _stream(self.device)))
                if iters.value < 0:
                    lz.enqueued()
                else:
                    lz.set(iters.value)
            return x.reshape(b.shape), lz
        with _on(self.device):
            solve = _lib.lib().efgp_cg_solve_hermitian if hermitian else _lib.lib().efgp_cg_solve
            check(solve(self._h, _ptr(w), float(sigmasq), int(mode), _ptr(mi), _ptr(bs), _ptr(x),
                        bs.shape[0], float(tol), -1 if max_iter is None else int(max_iter),
                        1 if early_stopping else 0, 1 if batched else 0, 1 if x0_zero else 0,
                        ctypes.byref(iters), _stream(self.device)))
        return x.reshape(b.shape), int(iters.value)

    def lanczos(self, q, steps, mode, ws, sigmasq, breakdown_tol=1e-12):
        """`steps` lockstep Lanczos steps of the fused operator `mode` from the unit-norm rows of q (P, M);
        returns (alphas, betas), each (P, steps) float64 on the device.  No host synchronisation."""
        _require_cuda(q, "q")
        qs = q.to(self.cdtype).reshape(-1, self.size).clone().contiguous()   # the library uses it as work space
        P = qs.shape[0]
        alphas = torch.empty((P, int(steps)), dtype=torch.float64, device=self.device)
        betas = torch.empty_like(alphas)
        w = self._ws(ws)
        with _on(self.device):
            check(_lib.lib().efgp_lanczos(self._h, _ptr(w), float(sigmasq), int(mode), _ptr(qs), P, int(steps),
                                          float(breakdown_tol), _ptr(alphas), _ptr(betas), _stream(self.device)))
        return alphas, betas


# ----------------------------------------------------------------------------- fused fit, bounding box, communicator
_BOX_RING = None
_BOX_NEXT = 0


def bbox_async(x, sharded=False):
    """Enqueue [max x_a .., -min x_a ..] of x (N, d) into pinned host memory (efgp_bbox_async; MAX-reduced over the
    library communicator when sharded).  Returns (pinned float64 tensor of 2 d entries, event): the caller does its
    L-independent host work, then event.synchronize() and reads the tensor."""
    global _BOX_RING, _BOX_NEXT
    _require_cuda(x, "x")
    if x.ndim == 1:
        x = x[:, None]
    x = x.contiguous()
    N, d = x.shape
    if _BOX_RING is None:
        _BOX_RING = torch.zeros((16, 8), dtype=torch.float64).pin_memory()
    buf = _BOX_RING[_BOX_NEXT]
    _BOX_NEXT = (_BOX_NEXT + 1) % 16
    with _on(x.device):
        check(_lib.lib().efgp_bbox_async(_ptr(x), int(N), int(d), _dtype_code(x.dtype), 1 if sharded else 0,
                                         ctypes.c_void_p(buf.data_ptr()), _stream(x.device)))
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(x.device))
    return buf[:2 * d], ev


BBOX_STATIC_MAX_POINTS = 134217728      # EFGP_BBOX_STATIC_MAX_POINTS of include/efgp_b200.h


def bbox_static(x, out):
    """efgp_bbox_static: [max x_a .., -min x_a ..] of x (N, d) into `out` (>= 2 d float64 on x's device, supplied by the
    caller) with the partials in the library's static device array: the call allocates nothing (see csrc/noalloc.cu;
    calls that overlap on the GPU race on that array).  x must already be contiguous -- no copy is made here."""
    _require_cuda(x, "x")
    _require_cuda(out, "out")
    if x.ndim == 1:
        x = x[:, None]
    if not x.is_contiguous():
        raise _lib.EfgpError("bbox_static: x must be contiguous (this entry point makes no copies)")
    N, d = x.shape
    if out.dtype != torch.float64 or out.numel() < 2 * d or not out.is_contiguous() or out.device != x.device:
        raise _lib.EfgpError("bbox_static: out must be a contiguous float64 tensor of >= 2 d entries on x's device")
    with _on(x.device):
        check(_lib.lib().efgp_bbox_static(_ptr(x), int(N), int(d), _dtype_code(x.dtype), _ptr(out), _stream(x.device)))
    return out


def fit(plan, y, args, ws, vr, minv, beta, lazy):
    """efgp_fit on the points of `plan` (a NufftPlan): fills ws, vr = [F^*1 ; ws F^*y], minv, beta (in/out) and returns
    the library's Toeplitz handle.  `lazy` is the LazyIters that receives the iteration counters."""
    handle = ctypes.c_void_p()
    iters = ctypes.c_int(0)
    with _on(plan.device):
        check(_lib.lib().efgp_fit(plan._h, _ptr(y), ctypes.byref(args), _ptr(ws), _ptr(vr), _ptr(minv), _ptr(beta),
                                  ctypes.byref(handle), lazy.ptr, ctypes.byref(iters), _stream(plan.device)))
        if iters.value < 0:
            lazy.enqueued()
        else:
            lazy.set(iters.value)
    plan.h = float(args.h)
    return handle


_COMM_KEY = None


def ensure_comm(group):
    """Create the library's NCCL communicator over the ranks of `group` (once per group): rank 0 draws the unique id,
    torch.distributed broadcasts its 128 bytes, every rank joins (efgp_comm_unique_id / efgp_comm_init)."""
    global _COMM_KEY
    import torch.distributed as dist
    key = (id(group), dist.get_world_size(group))
    if _COMM_KEY == key:
        return
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    idbuf = (ctypes.c_char * 128)()
    if rank == 0:
        check(_lib.lib().efgp_comm_unique_id(ctypes.cast(idbuf, ctypes.c_void_p)))
    dev = torch.device("cuda", torch.cuda.current_device())
    backend = dist.get_backend(group)
    t = torch.frombuffer(bytearray(idbuf.raw), dtype=torch.uint8).clone()
    if backend == "nccl":
        t = t.to(dev)
    src = dist.get_global_rank(group, 0) if hasattr(dist, "get_global_rank") else 0
    dist.broadcast(t, src=src, group=group)
    raw = bytes(t.cpu().numpy().tobytes())
    idbuf2 = ctypes.create_string_buffer(raw, 128)
    # NCCL prints its version banner to STDOUT at the first communicator of a process when NCCL_DEBUG is set; callers
    # (bench.py) own stdout, so the banner is sent to stderr for the duration of the call
    import os
    import sys
    sys.stdout.flush()
    saved = os.dup(1)
    try:
        os.dup2(2, 1)
        check(_lib.lib().efgp_comm_init(int(rank), int(world), ctypes.cast(idbuf2, ctypes.c_void_p)))
    finally:
        os.dup2(saved, 1)
        os.close(saved)
    _COMM_KEY = key


def comm_allreduce(t, op="sum"):
    """In-place all-reduce of a real device tensor over the library communicator (efgp_allreduce)."""
    _require_cuda(t, "t")
    assert t.is_contiguous() and t.dtype in (torch.float32, torch.float64)
    with _on(t.device):
        check(_lib.lib().efgp_allreduce(_ptr(t), int(t.numel()), _dtype_code(t.dtype), 0 if op == "sum" else 1,
                                        _stream(t.device)))
    return t


# ----------------------------------------------------------------------------- posterior variance stages
def xcorr_mean(us, etas, ws_real, mtot, d):
    """Lag sums of diag_sums_nd (efgp_xcorr_mean): us (J, M) complex, etas (J, M) real, ws_real (M,) or None ->
    ((2 mtot - 1),) * d complex, FFT ordering."""
    _require_cuda(us, "us")
    J, M = us.shape
    rdtype = real_of(us.dtype)
    us = us.contiguous()
    etas = etas.to(device=us.device, dtype=rdtype).contiguous()
    w = None if ws_real is None else ws_real.to(device=us.device, dtype=rdtype).reshape(-1).contiguous()
    out = torch.empty((2 * mtot - 1,) * d, dtype=us.dtype, device=us.device)
    with _on(us.device):
        check(_lib.lib().efgp_xcorr_mean(_ptr(us), _ptr(etas), _ptr(w), int(J), int(d), int(mtot), _dtype_code(rdtype),
                                         _ptr(out), _stream(us.device)))
    return out


def variance_rows(xb, h, mtot, ws_real):
    """rhs[i] = ws * conj(exp(2 pi i x_i . xi)) for the 'regular' variance (efgp_variance_rows): (b, mtot^d) complex."""
    _require_cuda(xb, "x")
    xb = xb.contiguous()
    b, d = xb.shape
    w = ws_real.to(device=xb.device, dtype=xb.dtype).reshape(-1).contiguous()
    rhs = torch.empty((b, mtot ** d), dtype=cmplx_of(xb.dtype), device=xb.device)
    with _on(xb.device):
        check(_lib.lib().efgp_variance_rows(_ptr(xb), int(b), int(d), int(mtot), float(h), _ptr(w),
                                            _dtype_code(xb.dtype), _ptr(rhs), _stream(xb.device)))
    return rhs


def variance_reduce(xb, h, mtot, ws_real, gamma):
    """var[i] = max(0, Re sum_k fx_i[k] ws[k] gamma[i, k]) (efgp_variance_reduce): (b,) real."""
    xb = xb.contiguous()
    b, d = xb.shape
    w = ws_real.to(device=xb.device, dtype=xb.dtype).reshape(-1).contiguous()
    g = gamma.to(cmplx_of(xb.dtype)).contiguous()
    var = torch.empty(b, dtype=xb.dtype, device=xb.device)
    with _on(xb.device):
        check(_lib.lib().efgp_variance_reduce(_ptr(xb), int(b), int(d), int(mtot), float(h), _ptr(w),
                                              _dtype_code(xb.dtype), _ptr(g), _ptr(var), _stream(xb.device)))
    return var
