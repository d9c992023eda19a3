"""ctypes binding of libefgp_b200.so (the C ABI in include/efgp_b200.h).

There is no CPU fallback: if the shared library is missing or a tensor is not on a CUDA device
the call raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libefgp_b200.so")

_lib = None

c_i64p = ctypes.POINTER(ctypes.c_int64)
c_dblp = ctypes.POINTER(ctypes.c_double)
c_intp = ctypes.POINTER(ctypes.c_int)
vp = ctypes.c_void_p

class FitArgs(ctypes.Structure):
    """efgp_fit_args of include/efgp_b200.h, field for field."""
    _fields_ = [("family", ctypes.c_int), ("c", ctypes.c_double), ("a", ctypes.c_double), ("b", ctypes.c_double),
                ("p", ctypes.c_double), ("h", ctypes.c_double), ("hd", ctypes.c_double), ("mtot", ctypes.c_int),
                ("sigmasq", ctypes.c_double), ("nufft_eps", ctypes.c_double), ("gen_eps", ctypes.c_double),
                ("cg_tol", ctypes.c_double), ("max_iter", ctypes.c_int), ("precondition", ctypes.c_int),
                ("force_pow2", ctypes.c_int), ("sharded", ctypes.c_int), ("warm_start", ctypes.c_int)]


_SIGNATURES = {
    "efgp_version": (ctypes.c_int, []),
    "efgp_last_error": (ctypes.c_char_p, []),
    "efgp_timing_enable": (ctypes.c_int, [ctypes.c_int]),
    "efgp_timing_report": (ctypes.c_int, [ctypes.c_char_p, ctypes.c_int64]),
    "efgp_pool_trim": (ctypes.c_int, []),
    "efgp_set_deterministic": (ctypes.c_int, [ctypes.c_int]),
    "efgp_measure_fp64_peak": (ctypes.c_int, [ctypes.POINTER(ctypes.c_double), ctypes.c_void_p]),
    "efgp_nufft_create": (ctypes.c_int, [ctypes.POINTER(vp), ctypes.c_int, ctypes.c_int]),
    "efgp_nufft_set_points": (ctypes.c_int, [vp, vp, ctypes.c_int64, ctypes.c_double, c_dblp, vp]),
    "efgp_nufft_destroy": (None, [vp]),
    "efgp_nufft_type1": (ctypes.c_int, [vp, vp, ctypes.c_int, ctypes.c_int, c_i64p, ctypes.c_double,
                                        ctypes.c_int, ctypes.c_int, vp, vp, vp]),
    "efgp_nufft_type1_pair": (ctypes.c_int, [vp, vp, c_i64p, ctypes.c_double, c_i64p, ctypes.c_double, vp, vp, vp, vp]),
    "efgp_nufft_type2": (ctypes.c_int, [vp, vp, ctypes.c_int, c_i64p, ctypes.c_double, ctypes.c_int,
                                        ctypes.c_int, vp, ctypes.c_int, vp, vp]),
    "efgp_nufft_plan_info": (ctypes.c_int, [vp, c_i64p, ctypes.c_double, c_i64p, c_intp]),
    "efgp_toeplitz_create": (ctypes.c_int, [ctypes.POINTER(vp), vp, ctypes.c_int, c_i64p, ctypes.c_int,
                                            ctypes.c_int, vp]),
    "efgp_toeplitz_destroy": (None, [vp]),
    "efgp_toeplitz_fft_shape": (ctypes.c_int, [vp, c_i64p]),
    "efgp_toeplitz_apply": (ctypes.c_int, [vp, vp, ctypes.c_int, vp, ctypes.c_double, ctypes.c_int, vp, vp]),
    "efgp_cg_solve": (ctypes.c_int, [vp, vp, ctypes.c_double, ctypes.c_int, vp, vp, vp, ctypes.c_int,
                                     ctypes.c_double, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, c_intp,
                                     vp]),
    "efgp_cg_solve_hermitian": (ctypes.c_int, [vp, vp, ctypes.c_double, ctypes.c_int, vp, vp, vp, ctypes.c_int,
                                               ctypes.c_double, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                               c_intp, vp]),
    "efgp_cg_solve_async": (ctypes.c_int, [vp, vp, ctypes.c_double, ctypes.c_int, vp, vp, vp, ctypes.c_int,
                                           ctypes.c_double, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                           ctypes.c_int, vp, c_intp, vp]),
    "efgp_lanczos": (ctypes.c_int, [vp, vp, ctypes.c_double, ctypes.c_int, vp, ctypes.c_int, ctypes.c_int,
                                    ctypes.c_double, vp, vp, vp]),
    "efgp_nufft_rescale": (ctypes.c_int, [vp, ctypes.c_double, c_dblp]),
    "efgp_nufft_carry_index": (ctypes.c_int, [vp, ctypes.c_int]),
    "efgp_comm_unique_id": (ctypes.c_int, [vp]),
    "efgp_comm_init": (ctypes.c_int, [ctypes.c_int, ctypes.c_int, vp]),
    "efgp_comm_destroy": (ctypes.c_int, []),
    "efgp_comm_world": (ctypes.c_int, []),
    "efgp_allreduce": (ctypes.c_int, [vp, ctypes.c_int64, ctypes.c_int, ctypes.c_int, vp]),
    "efgp_bbox_async": (ctypes.c_int, [vp, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_int, vp, vp]),
    "efgp_bbox_static": (ctypes.c_int, [vp, ctypes.c_int64, ctypes.c_int, ctypes.c_int, vp, vp]),
    "efgp_xcorr_mean": (ctypes.c_int, [vp, vp, vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, vp, vp]),
    "efgp_variance_rows": (ctypes.c_int, [vp, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_double, vp, ctypes.c_int,
                                          vp, vp]),
    "efgp_variance_reduce": (ctypes.c_int, [vp, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_double, vp,
                                            ctypes.c_int, vp, vp, vp]),
    "efgp_nufft_type1_probes": (ctypes.c_int, [vp, ctypes.c_uint64, ctypes.c_int, ctypes.c_int, c_i64p, ctypes.c_double, vp, vp,
                                               ctypes.POINTER(ctypes.c_uint64), vp]),
    "efgp_nufft_type2_probe_dots": (ctypes.c_int, [vp, ctypes.c_uint64, ctypes.c_int, ctypes.c_int, vp, c_i64p,
                                                   ctypes.c_double, vp, vp, ctypes.POINTER(ctypes.c_uint64), vp]),
    "efgp_nufft_probes_export": (ctypes.c_int, [vp, ctypes.c_uint64, ctypes.c_int, ctypes.c_int, c_i64p, ctypes.c_double, vp,
                                                ctypes.POINTER(ctypes.c_uint64), vp]),
    "efgp_fit": (ctypes.c_int, [vp, vp, ctypes.POINTER(FitArgs), vp, vp, vp, vp, ctypes.POINTER(vp), vp, c_intp, vp]),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)


class EfgpError(RuntimeError):
    pass


def lib():
    """Load (once) and return the shared library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise EfgpError(
                f"{LIB_PATH} not found: build it with `python gp-quadrature_b200/build.py` "
                "(there is no CPU or PyTorch fallback for the EFGP hot path)")
        l = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(code):
    if code != 0:
        msg = lib().efgp_last_error()
        raise EfgpError(f"libefgp_b200 error {code}: {msg.decode() if msg else '?'}")


def i64_array(vals):
    vals = [int(v) for v in vals]
    return (ctypes.c_int64 * len(vals))(*vals)


def dbl_array(vals):
    vals = [float(v) for v in vals]
    return (ctypes.c_double * len(vals))(*vals)


def set_deterministic(on=True):
    """Bit-reproducible type-1 transforms (fixed-point accumulation): see efgp_set_deterministic in the header."""
    check(lib().efgp_set_deterministic(1 if on else 0))


def timing_enable(on=True):
    """Start (and reset) / stop per-stage device timing inside the library."""
    check(lib().efgp_timing_enable(1 if on else 0))


def timing_report():
    """{stage: {"scopes": n, "launches": n, "units": points processed, "ms": total}} since timing_enable(True); synchronises."""
    buf = ctypes.create_string_buffer(1 << 16)
    check(lib().efgp_timing_report(buf, len(buf)))
    out = {}
    for line in buf.value.decode().splitlines():
        name, scopes, launches, units, ms = line.rsplit(",", 4)
        out[name] = {"scopes": int(scopes), "launches": int(launches), "units": int(units), "ms": float(ms)}
    return out


def measure_fp64_peak(stream=None):
    """Measured FP64 FMA throughput of the current device in TFLOP/s (diagnostic for bench.py)."""
    out = ctypes.c_double(0.0)
    check(lib().efgp_measure_fp64_peak(ctypes.byref(out), ctypes.c_void_p(stream or 0)))
    return float(out.value)
