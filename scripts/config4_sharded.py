"""BASELINE config 4 (3-D SE hyperparameter-gradient step, N = 1e7, 64 Hutchinson probes) with the probes sharded
over the ranks (opts["shard_probes"]): every rank holds all points, runs T/G probes, ONE all-reduce of the partial
trace sums.  Launch: python -m torch.distributed.run --nproc-per-node G scripts/config4_sharded.py  (G = 1 works too).
Prints one JSON line on rank 0; time = max over ranks of the CUDA-event time of the second (warm) step."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gp-quadrature_b200")); sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
import efgpnd
from kernels import SquaredExponential

world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
if "MASTER_ADDR" not in os.environ:
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT="29533", RANK="0", WORLD_SIZE="1")
dist.init_process_group("nccl", device_id=dev)
N, d, T = int(float(os.environ.get("C4_N", "1e7"))), 3, int(os.environ.get("C4_T", "64"))
g = torch.Generator(device=dev).manual_seed(0)                 # same points on every rank
x = torch.rand(N, d, dtype=torch.float64, device=dev, generator=g) * 2 - 1
y = (torch.sin(3 * x[:, 0]) * torch.cos(4 * x[:, 1]) + 0.3 * torch.sin(5 * x[:, 2])
     + 0.2 * torch.randn(N, dtype=torch.float64, device=dev, generator=g))
k = SquaredExponential(dimension=d, init_lengthscale=0.1, init_variance=1.0)
m = efgpnd.EFGPND(x, y, k, sigmasq=0.04, eps=1e-4, estimate_params=False, opts={"shard_probes": True})
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
times, grads = [], None
for rep in range(3):
    torch.manual_seed(1234)                                    # identical probe seeds on every rank and repeat
    dist.barrier(); torch.cuda.synchronize()
    e0.record()
    grads = m.compute_gradients(trace_samples=T, cg_tol=1e-5, nufft_eps=1e-5, apply_gradients=False)
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    times.append(float(t.item()))
st = m.last_gradient_stats
if rank == 0:
    print(json.dumps({"config": 4, "n_gpus": world, "N": N, "d": d, "T": T, "mtot": st["mtot"],
                      "first_step_ms": times[0], "step_ms": min(times[1:]), "probe_points_per_s": N * T / min(times[1:]) * 1e3,
                      "trace_cg_iters": st["trace_cg_iters"], "grad": [float(v) for v in grads],
                      "mem_GB": torch.cuda.max_memory_allocated() / 1e9}), flush=True)
dist.destroy_process_group()
