"""World-size-2 gloo test (CPU) of the host-side multi-GPU logic: point sharding, the bounding-box MAX
all-reduce that fixes h, and the single SUM all-reduce of the packed mode grids.  The per-rank transforms
are done by the exact-NUDFT oracle here (the CUDA kernels need a GPU; tests/test_ops_gpu.py checks their
linearity under sharding), so this verifies that the sharded fit assembles exactly the unsharded system."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, PKG


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    import sys
    for p in (PKG, ROOT, os.path.join(ROOT, "oracle")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import efgpnd
    from pytorch_finufft import functional as nudft
    torch.manual_seed(7)                      # same full data set on every rank; each takes its shard
    N, d, h, m = 1001, 2, 0.31, 4
    x = torch.rand(N, d, dtype=torch.float64) * 3 - 1
    y = torch.randn(N, dtype=torch.float64)
    lo, hi = efgpnd.shard_bounds(N, world, rank)
    xs, ys = x[lo:hi], y[lo:hi]
    group = efgpnd._point_shard_group({"shard_points": True})
    assert group is not None
    L = efgpnd._domain_length(xs, group)
    phi = (2 * np.pi * h * xs).T.contiguous()
    mt = 2 * m + 1
    v = nudft.finufft_type1(phi, torch.ones(hi - lo, dtype=torch.complex128), (4 * m + 1,) * d, isign=-1)
    ws = torch.linspace(0.5, 1.5, mt ** d, dtype=torch.float64)
    rhs = ws * nudft.finufft_type1(phi, ys.to(torch.complex128), (mt,) * d, isign=-1).reshape(-1)
    v_all, rhs_all = efgpnd._sum_over_shards(v, rhs, group)
    torch.save({"L": L, "v": v_all, "rhs": rhs_all, "n": hi - lo}, os.path.join(out_dir, f"r{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_point_sharded_fit_assembles_the_unsharded_system(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    import sys
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    from pytorch_finufft import functional as nudft
    import efgpnd
    torch.manual_seed(7)
    N, d, h, m = 1001, 2, 0.31, 4
    x = torch.rand(N, d, dtype=torch.float64) * 3 - 1
    y = torch.randn(N, dtype=torch.float64)
    phi = (2 * np.pi * h * x).T.contiguous()
    mt = 2 * m + 1
    v = nudft.finufft_type1(phi, torch.ones(N, dtype=torch.complex128), (4 * m + 1,) * d, isign=-1)
    ws = torch.linspace(0.5, 1.5, mt ** d, dtype=torch.float64)
    rhs = ws * nudft.finufft_type1(phi, y.to(torch.complex128), (mt,) * d, isign=-1).reshape(-1)
    L = efgpnd._domain_length(x, None)
    outs = [torch.load(os.path.join(str(tmp_path), f"r{r}.pt")) for r in range(world)]
    assert sum(o["n"] for o in outs) == N
    for o in outs:
        assert o["L"] == L                                   # bit-identical h on every rank
        assert float((o["v"] - v).abs().max() / v.abs().max()) < 1e-13
        assert float((o["rhs"] - rhs).abs().max() / rhs.abs().max()) < 1e-13
    assert torch.equal(outs[0]["v"], outs[1]["v"]) and torch.equal(outs[0]["rhs"], outs[1]["rhs"])  # replicas agree


def test_shard_bounds_cover_all_points():
    import efgpnd
    for n, w in [(10, 3), (100000001, 8), (7, 8), (0, 2)]:
        b = [efgpnd.shard_bounds(n, w, r) for r in range(w)]
        assert b[0][0] == 0 and b[-1][1] == n
        assert all(b[r][1] == b[r + 1][0] for r in range(w - 1))
        assert max(hi - lo for lo, hi in b) - min(hi - lo for lo, hi in b) <= 1


def test_shard_points_requires_process_group():
    import pytest
    import efgpnd
    from efgp_b200._lib import EfgpError
    assert efgpnd._point_shard_group({}) is None
    if not dist.is_initialized():
        with pytest.raises(EfgpError):
            efgpnd._point_shard_group({"shard_points": True})
