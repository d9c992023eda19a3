"""Probe-sharded gradient (north_star: Hutchinson probes sharded per GPU, scalar all-reduce): two ranks
sharing the one test GPU (gloo collectives) must reproduce the single-rank result of the same seeded
probes, and the iteration statistics."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, PKG

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    import sys
    for p in (PKG, ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import efgpnd
    from kernels import SquaredExponential
    torch.manual_seed(5)
    N, d = 4000, 2
    x = (torch.rand(N, d, dtype=torch.float64) * 2 - 1).cuda()
    y = (torch.sin(3 * x[:, 0]) * torch.cos(2 * x[:, 1]) + 0.1 * torch.randn(N, dtype=torch.float64, device="cuda"))
    kern = SquaredExponential(dimension=d, init_lengthscale=0.3, init_variance=1.0)
    m = efgpnd.EFGPND(x, y, kern, sigmasq=1.0, eps=1e-4, opts={"shard_probes": True}, estimate_params=False)
    torch.manual_seed(1234)
    g = m.compute_gradients(trace_samples=7, cg_tol=1e-9, nufft_eps=1e-9, apply_gradients=False)
    torch.save({"grad": g.detach().cpu(), "stats": {k: v for k, v in m.last_gradient_stats.items()}},
               os.path.join(out_dir, f"w{world}_r{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_reproduce_one_rank(tmp_path):
    for world in (1, 2):
        mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    one = torch.load(os.path.join(str(tmp_path), "w1_r0.pt"))
    for r in range(2):
        two = torch.load(os.path.join(str(tmp_path), f"w2_r{r}.pt"))
        rel = float((two["grad"] - one["grad"]).norm() / one["grad"].norm())
        assert rel < 1e-7, rel          # bar: 1e-5 (BASELINE.json); CG tolerance of the run is 1e-9
        assert two["stats"]["trace_num_rhs"] == one["stats"]["trace_num_rhs"] == 2 * 7
        assert two["stats"]["trace_samples"] == 7
        # iteration counts are statistics, not results: the fp64 atomics of the spread make the right-hand sides
        # differ in the last bits from run to run, and a row that sits at the 1e-9 tolerance after ~460
        # iterations may need a few more (460 and 474 both occur): allow 5 %
        it1, it2 = one["stats"]["trace_cg_iters"], two["stats"]["trace_cg_iters"]
        assert abs(it2 - it1) <= max(2, 0.05 * it1), (it1, it2)
