"""efgp_bbox_static (csrc/noalloc.cu): the bounding-box reduction with its scratch in a file-scope static __device__
array, for callers that cannot call cudaMalloc.  Called through the C ABI (ctypes on libefgp_b200.so) and compared with
the pooled entry point efgp_bbox_async on the same points -- max / min are exact, so the bar is bit equality, tighter
than any tolerance the other tests use -- on two calls in a row (the second one smaller, so the array still holds the
first call's partials in the slots the second call must not read).

This file sorts last on purpose: a failure here must not hide another test.
It passes alone in a fresh process and after the rest of the suite (it keeps no state besides the library itself).
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_POINTS = 134217728            # EFGP_BBOX_STATIC_MAX_POINTS of include/efgp_b200.h


def _c_abi():
    from efgp_b200 import _lib
    return _lib.lib(), _lib


def _bbox_pooled(lib, x):
    """The existing entry point: result through pinned host memory."""
    N, d = x.shape
    pinned = torch.full((8,), float("nan"), dtype=torch.float64).pin_memory()
    st = torch.cuda.current_stream(x.device).cuda_stream
    rc = lib.efgp_bbox_async(ctypes.c_void_p(x.data_ptr()), N, d, 0 if x.dtype == torch.float32 else 1, 0,
                             ctypes.c_void_p(pinned.data_ptr()), ctypes.c_void_p(st))
    assert rc == 0, lib.efgp_last_error()
    torch.cuda.synchronize()
    return pinned[:2 * d].clone()


def _bbox_static(lib, x, out):
    N, d = x.shape
    st = torch.cuda.current_stream(x.device).cuda_stream
    rc = lib.efgp_bbox_static(ctypes.c_void_p(x.data_ptr()), N, d, 0 if x.dtype == torch.float32 else 1,
                              ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(st))
    assert rc == 0, lib.efgp_last_error()
    torch.cuda.synchronize()
    return out[:2 * d].cpu()


def _points(N, d, dt, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(N, d, dtype=torch.float64, device="cuda", generator=g) * (1.0 + seed) + 0.25 * seed
    return x.to(dt).contiguous()


@pytest.mark.parametrize("dt", [torch.float64, torch.float32])
@pytest.mark.parametrize("d", [1, 2, 3])
def test_static_scratch_bbox_matches_pooled_bbox_twice_in_a_row(d, dt):
    lib, _ = _c_abi()
    out = torch.full((8,), float("nan"), dtype=torch.float64, device="cuda")
    # first call: many tiles, ragged tail; second call: fewer tiles and different data (stale partials stay behind);
    # then the edge sizes around one 8192-point tile
    for k, N in enumerate([1_000_003, 20_001, 8193, 8192, 8191, 1]):
        x = _points(N, d, dt, seed=k)
        want = _bbox_pooled(lib, x)
        out.fill_(float("nan"))
        got = _bbox_static(lib, x, out)
        assert torch.equal(got, want), (N, d, dt, got, want)
        # and against torch itself
        ref = torch.cat([x.max(0).values, -x.min(0).values]).double().cpu()
        assert torch.equal(got, ref)
        assert bool(torch.isnan(out[2 * d:]).all())        # nothing written past the 2 d results


def test_static_scratch_bbox_back_to_back_without_host_sync():
    """Two calls enqueued on one stream with no synchronisation between them (stream order serialises them)."""
    lib, _ = _c_abi()
    xa, xb = _points(3_000_000, 2, torch.float64, 11), _points(500_000, 2, torch.float64, 12)
    oa = torch.zeros(4, dtype=torch.float64, device="cuda")
    ob = torch.zeros(4, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for _ in range(2):
        assert lib.efgp_bbox_static(ctypes.c_void_p(xa.data_ptr()), xa.shape[0], 2, 1, ctypes.c_void_p(oa.data_ptr()), st) == 0
        assert lib.efgp_bbox_static(ctypes.c_void_p(xb.data_ptr()), xb.shape[0], 2, 1, ctypes.c_void_p(ob.data_ptr()), st) == 0
    torch.cuda.synchronize()
    assert torch.equal(oa.cpu(), _bbox_pooled(lib, xa))
    assert torch.equal(ob.cpu(), _bbox_pooled(lib, xb))


def test_static_scratch_bbox_takes_no_device_memory():
    """Free device memory is the same before and after calls (the array is part of the loaded module)."""
    lib, _ = _c_abi()
    x = _points(2_000_000, 3, torch.float32, 5)
    out = torch.zeros(8, dtype=torch.float64, device="cuda")
    _bbox_static(lib, x, out)                      # module load / symbol lookup happen here at the latest
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        _bbox_static(lib, x, out)
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0


def test_static_scratch_bbox_rejects_sizes_beyond_the_array():
    lib, _lib = _c_abi()
    x = _points(16, 2, torch.float64, 1)            # never dereferenced: the size check comes first
    out = torch.full((4,), 7.0, dtype=torch.float64, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.efgp_bbox_static(ctypes.c_void_p(x.data_ptr()), MAX_POINTS + 1, 2, 1, ctypes.c_void_p(out.data_ptr()), st)
    assert rc == 1
    msg = lib.efgp_last_error().decode()
    assert "maximum" in msg and str(MAX_POINTS) in msg
    rc = lib.efgp_bbox_static(ctypes.c_void_p(x.data_ptr()), 16, 4, 1, ctypes.c_void_p(out.data_ptr()), st)
    assert rc == 1 and "dim" in lib.efgp_last_error().decode()
    torch.cuda.synchronize()
    assert bool((out == 7.0).all())                 # a refused call writes nothing
    # the host-side wrapper raises the library's error class
    from efgp_b200 import ops
    assert ops.BBOX_STATIC_MAX_POINTS == MAX_POINTS
    with pytest.raises(_lib.EfgpError):
        ops.bbox_static(x.cpu(), out)
    got = ops.bbox_static(x, out)
    torch.cuda.synchronize()
    assert torch.equal(got.cpu(), _bbox_pooled(lib, x))
