#!/usr/bin/env python
"""Headline benchmark: 2-D squared-exponential EFGP posterior-mean solve on N = 1e8 synthetic points
(BASELINE.json configs[2]), eps = 1e-4, fp64.

One step = the whole hot path on one batch of points: bin-sort of x, the F^*1 Toeplitz generator
(type-1 at eps 6e-8), ws * F^* y (type-1), Toeplitz spectrum, Jacobi-preconditioned CG for beta, and
the posterior mean Re F(ws beta) at the N training points (type-2)  --  exactly
EFGPND(x, y, ...).predict(x, return_variance=False) of the reference (efgpnd.py:710-822, 918-922).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--npts N]

Multi-GPU (torchrun, one rank per GPU): the N points are sharded over the ranks ("strong" scaling:
total N fixed), each rank spreads its shard, ONE NCCL all-reduce sums the two mode grids, every
rank solves the same M x M system and interpolates at its own points.

`--impl reference` times the reference's CPU path (the oracle's restatement of efgpnd.py with the
C/OpenMP NUFFT of oracle/cnufft.c standing in for FINUFFT, which is not installable here) on a
bounded sample of the same workload, on all host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "gp-quadrature_b200")
for _p in (PKG, ROOT):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "efgp_posterior_mean_solve_throughput"
UNIT = "Mpts/s"
ELL, VAR, SIG2, EPS, NUFFT_EPS, CG_TOL = 0.05, 1.0, 0.1, 1e-4, 1e-4, 1e-4


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--npts", type=float, default=1e8, help="total number of points (all ranks)")
    ap.add_argument("--cpu-sample", type=float, default=4e6, help="points in the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def synth(n, d, seed, device, dtype):
    """x ~ U[-1,1]^d, y = analytic field + N(0, 0.2^2)  (test_timing_profiling.py:25-44 recipe)."""
    import torch
    g = torch.Generator(device=device).manual_seed(seed)
    x = torch.rand(n, d, dtype=dtype, device=device, generator=g) * 2 - 1
    r2 = (x * x).sum(1)
    y = (torch.sin(3 * x[:, 0]) * torch.cos(4 * x[:, 1])
         + 0.5 * torch.exp(-((x[:, 0] - 0.3) ** 2 + (x[:, 1] + 0.3) ** 2) / 0.3)
         + 0.7 * torch.sin(2 * torch.pi * r2))
    y = y + 0.2 * torch.randn(n, dtype=dtype, device=device, generator=g)
    return x, y


# ----------------------------------------------------------------------------- CPU reference arm
def _host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_solve_step(x, y):
    """One posterior-mean solve with the oracle's restatement of the reference + the fast CPU NUFFT."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch
    import efgp_oracle as O
    import fast_nufft as F
    # all host cores, also under torchrun (which exports OMP_NUM_THREADS=1)
    F.NTHREADS = _host_threads()
    torch.set_num_threads(_host_threads())
    kern = O.OracleKernel("se", 2, ELL, VAR)
    f = O.fit(x, y, kern, SIG2, EPS, NUFFT_EPS, cg_tol=CG_TOL, backend=F)
    mean = O.predict_mean(f, x, NUFFT_EPS, backend=F)
    return f, mean


def cpu_baseline(n_sample, reps):
    import torch
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import fast_nufft as F
    x, y = synth(int(n_sample), 2, 1234, "cpu", torch.float64)
    cpu_solve_step(x[:20000], y[:20000])  # warm the libraries
    best = float("inf")
    for _ in range(reps):
        t0 = time.perf_counter()
        f, _ = cpu_solve_step(x, y)
        best = min(best, time.perf_counter() - t0)
    cores = _host_threads()
    return {"value": n_sample / best / 1e6, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{int(n_sample)} of the workload's points, full solve (2 type-1, CG {f.iters} its, 1 type-2), "
                      f"best of {reps}; oracle restatement of efgpnd.py + C/OpenMP ES-kernel NUFFT (FINUFFT absent)",
            "seconds": best, "os_cpu_count": os.cpu_count()}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    reps = max(1, args.steps)
    n_s = args.cpu_sample
    import torch
    x, y = synth(int(n_s), 2, 1234, "cpu", torch.float64)
    for _ in range(min(args.warmup, 1)):
        cpu_solve_step(x[:200000], y[:200000])
    t0 = time.perf_counter()
    for _ in range(reps):
        f, _ = cpu_solve_step(x, y)
    dt = (time.perf_counter() - t0) / reps
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import fast_nufft as F
    v = n_s / dt / 1e6
    cores = _host_threads()
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": reps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, 1, f.mtot),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{int(n_s)} points per step (bounded sample of the N={int(args.npts)} workload)",
                             "os_cpu_count": os.cpu_count()},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(args, world, mtot=None):
    return {"workload": "BASELINE.json configs[2]: 2-D squared-exponential EFGP posterior-mean solve "
                        "(sort + F^*1 @6e-8 + F^*y + Toeplitz + PCG + mean at the N training points)",
            "N_total": int(args.npts), "d": 2, "lengthscale": ELL, "variance": VAR, "sigmasq": SIG2, "eps": EPS,
            "nufft_eps": NUFFT_EPS, "cg_tolerance": CG_TOL, "mtot": mtot,
            "parallelism": f"points sharded over {world} GPU(s), NCCL all-reduce of the mode grids" if world > 1
            else "single GPU",
            "l2_policy": "inputs (>= 2.4 GB per step) are larger than the 126 MB L2; the sort is redone every step"}


# ----------------------------------------------------------------------------- clocks
class NvmlClockSampler:
    """SM clock / throttle reasons sampled through NVML every 5 ms from a thread: the timed region of a sharded run is
    only tens of milliseconds, shorter than the start-up of an `nvidia-smi -lms` process (ClockSampler below, kept as
    the fallback when pynvml is missing)."""
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, gpu_index):
        import threading
        import pynvml
        pynvml.nvmlInit()
        self.nv = pynvml
        self.h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        self.sm, self.pw, self.reasons = [], [], set()
        self._stop = threading.Event()
        self._sample()
        self.sm.clear()                   # the probe sample above is outside the timed region
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()

    def _sample(self):
        nv = self.nv
        self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
        try:
            self.pw.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
        except Exception:
            pass
        try:
            get = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
            mask = int(get(self.h))
            for bit, name in self.REASONS:
                if mask & bit:
                    self.reasons.add(name)
        except Exception:
            pass

    def _run(self):
        while not self._stop.is_set():
            self._sample()
            self._stop.wait(0.005)

    def stop(self):
        self._stop.set()
        self.t.join(timeout=2)
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.mx, "reasons": ["no samples"]}
        hi = sorted(self.sm)[len(self.sm) // 2:]            # samples under load = the upper half
        return {"sm_mhz": sorted(hi)[len(hi) // 2], "sm_max_mhz": self.mx, "reasons": sorted(self.reasons),
                "samples": len(self.sm), "power_w_max": max(self.pw) if self.pw else None, "source": "nvml"}


def make_clock_sampler(gpu_index):
    try:
        return NvmlClockSampler(gpu_index)
    except Exception:
        return ClockSampler(gpu_index)


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={gpu_index}", f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons, pw = [], [], set(), []
        for line in self.f.read().splitlines():
            c = [t.strip() for t in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1]))
                mx.append(float(c[2]))
                pw.append(float(c[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        hi = sorted(sm)[len(sm) // 2:]            # samples under load = the upper half
        return {"sm_mhz": sorted(hi)[len(hi) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(pw)}


# ----------------------------------------------------------------------------- algorithmic bytes
def algorithmic_bytes(stage, n, d=2, word=8):
    """Bytes `stage` must move to process n points."""
    if stage.startswith("spread"):
        return n * word * (d + (0 if stage.endswith("_ones") else 1))      # coordinates (+ real strength)
    if stage.startswith("interp"):
        return n * (word * d + word)                                        # coordinates + real result
    per_point = {
        "sort.bucket_hist_kernel": word * d,                                # read x
        "sort.bucket_scatter_kernel": word * d + word * d + 2 + 2,          # read x; write folded coords, bin, A (uint16)
        "sort.bucket_scatter_kernel_pay": word * d + word + word * d + word + 2 + 2,   # ... plus the y payload in and out
        "sort.bin_scatter_kernel_pay": word * d + word + 2 + word * d + word + 4 + 4,  # ... plus y into the record
        "sort.bin_hist_kernel": 2,                                          # read level-2 bins
        "sort.bin_scatter_kernel": word * d + 2 + word * d + 4 + 4,         # read coords, bin; write record (coords, binv), B
        "permute.to_bucket_order": word + 2 + word,                         # read c, A; write tmp
        "permute.to_slot_order": 4 + word + word,                           # read Binv, tmp; write cs
        "unpermute.slot_to_bucket_order": 4 + word + word,                  # read B, r; write tmp
        "unpermute.from_bucket_order": 2 + word + word,                     # read A, tmp; write out
    }
    if stage in per_point:
        return n * per_point[stage]
    return None


def algorithmic_flops(stage, n, d=2):
    """FP64 flops (FMA = 2) the spreading / interpolation kernels need for n points: per dimension the even/odd
    Horner evaluation of the W piecewise polynomials of degree W + 2 (spread.cu es_eval), then W^d FMAs."""
    import re
    m = re.match(r"(spread|interp)\dd_w(\d+)_", stage)
    if not m:
        return None
    W = int(m.group(2))
    deg = W + 2
    ne, no = deg // 2 + 1, (deg + 1) // 2
    horner = (W // 2) * (2 * (ne + no - 2) + 3) + (2 * (ne - 1) if W % 2 else 0) + 4
    return n * (d * horner + 2 * W ** d + W)


def bind_to_gpu_numa_node(local):
    """Pin this rank to the CPUs next to its GPU so that its pinned staging buffers (first touch) and the
    H2D / D2H DMA stay on the GPU's own NUMA node; 8 ranks otherwise share one node's memory controllers."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(local)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bdf}/local_cpulist") as f:
            spec = f.read().strip()
        cpus = set()
        for part in spec.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return spec
    except Exception:
        pass
    return None


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback for the product path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bind_to_gpu_numa_node(local) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import efgpnd
    from efgp_b200 import _lib
    from kernels.squared_exponential import SquaredExponential

    n_total = int(args.npts)
    lo_pt, hi_pt = efgpnd.shard_bounds(n_total, world, rank)
    n_loc = hi_pt - lo_pt
    x, y = synth(n_loc, 2, 1000 + rank, dev, torch.float64)
    opts = {"cg_tolerance": CG_TOL, "shard_points": world > 1}
    stats = {}

    def solve(xd, yd):
        efgpnd._PLAN_CACHE.clear()                 # a step starts from raw points: sort included
        kern = SquaredExponential(dimension=2, init_lengthscale=ELL, init_variance=VAR)
        model = efgpnd.EFGPND(xd, yd, kern, sigmasq=SIG2, eps=EPS, nufft_eps=NUFFT_EPS, opts=opts,
                              estimate_params=False)
        mean, _ = model.predict(xd, return_variance=False)
        stats.update(model.last_fit_stats)
        return mean

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        mean = solve(x, y)
    barrier()

    # ---- device-resident timing (value) with per-stage timing inside the library
    sampler = make_clock_sampler(local) if rank == 0 else None
    _lib.timing_enable(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        mean = solve(x, y)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1) / args.steps
    stages = _lib.timing_report()
    _lib.timing_enable(False)
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    check = float(mean.double().square().mean().sqrt().item())

    # ---- end to end: pinned host buffers -> device -> solve -> result back on the host, every step
    e2e = None
    if not args.no_e2e:
        xh = torch.empty((n_loc, 2), dtype=torch.float64, pin_memory=True)
        yh = torch.empty((n_loc,), dtype=torch.float64, pin_memory=True)
        mh = torch.empty((n_loc,), dtype=torch.float64, pin_memory=True)
        xh.copy_(x)
        yh.copy_(y)
        xd, yd = torch.empty_like(x), torch.empty_like(y)

        def e2e_step():
            xd.copy_(xh, non_blocking=True)
            yd.copy_(yh, non_blocking=True)
            m = solve(xd, yd)
            mh.copy_(m, non_blocking=True)

        def timed(run_steps):
            barrier()
            e0.record()
            run_steps()
            e1.record()
            barrier()
            tt = torch.tensor([e0.elapsed_time(e1) / args.steps], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            return float(tt.item())

        # (1) serial: copy in, solve, copy out, one step after the other on one stream
        e2e_step()
        ms_serial = timed(lambda: [e2e_step() for _ in range(args.steps)])

        # (2) double-buffered, what a serving loop does: the upload of step k+1 (copy stream) runs while step k
        # is solved (compute stream) and the result of step k goes back on a third stream.  Every step still
        # uploads its own 24 B/pt and downloads its own 8 B/pt inside the timed region.
        cur = torch.cuda.current_stream()
        s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
        xdb, ydb = [xd, torch.empty_like(x)], [yd, torch.empty_like(y)]
        mhb = [mh, torch.empty((n_loc,), dtype=torch.float64, pin_memory=True)]
        ev_in = [torch.cuda.Event(), torch.cuda.Event()]
        ev_free = [torch.cuda.Event(), torch.cuda.Event()]

        def upload(b):
            with torch.cuda.stream(s_in):
                s_in.wait_event(ev_free[b])          # the solve that last read buffer b has finished
                xdb[b].copy_(xh, non_blocking=True)
                ydb[b].copy_(yh, non_blocking=True)
                ev_in[b].record(s_in)

        def pipelined(nsteps):
            s_in.wait_stream(cur)
            s_out.wait_stream(cur)
            upload(0)
            for k in range(nsteps):
                b = k & 1
                if k + 1 < nsteps:
                    upload(b ^ 1)                    # enqueued BEFORE solve(k): overlaps it
                cur.wait_event(ev_in[b])
                m = solve(xdb[b], ydb[b])
                ev_free[b].record(cur)
                with torch.cuda.stream(s_out):
                    s_out.wait_event(ev_free[b])
                    mhb[b].copy_(m, non_blocking=True)
                    m.record_stream(s_out)
            cur.wait_stream(s_in)
            cur.wait_stream(s_out)

        pipelined(2)
        ms_e2e = timed(lambda: pipelined(args.steps))
        e2e = {"value": n_total / ms_e2e / 1e3, "unit": UNIT, "ms_per_step": ms_e2e,
               "h2d_bytes_per_step": (xh.numel() + yh.numel()) * 8 * world,
               "d2h_bytes_per_step": mh.numel() * 8 * world,
               "mode": "double-buffered: H2D of step k+1 and D2H of step k overlap the solve of step k (3 streams); "
                       "every step uploads its own inputs and downloads its own result inside the timed region",
               "serial": {"value": n_total / ms_serial / 1e3, "ms_per_step": ms_serial,
                          "mode": "copy in, solve, copy out back to back on one stream"},
               "api": "EFGPND(x, y, SquaredExponential, ...).predict(x, return_variance=False) on buffers copied "
                      "from pinned host memory; posterior mean copied back"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant hand-written kernel (live CUDA-event durations from the library)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    own = {k: v for k, v in stages.items() if algorithmic_bytes(k, 1) is not None and v["launches"] > 0}
    per_stage = {}
    for k, v in stages.items():
        ent = {"ms_per_step": v["ms"] / args.steps, "launches_per_step": v["launches"] / args.steps}
        if k in own and v["ms"] > 0:
            # algorithmic bytes of all the points the stage processed / the CUDA-event time of its launches
            ent["achieved_gbs"] = algorithmic_bytes(k, v["units"]) / v["ms"] / 1e6
            ent["hbm_frac"] = ent["achieved_gbs"] / peak
        per_stage[k] = ent
    roofline = None
    if own:
        top = max(own, key=lambda k: own[k]["ms"])
        v = own[top]
        ach = algorithmic_bytes(top, v["units"]) / v["ms"] / 1e6
        # DRAM bytes per launch from the committed `ncu --set full` capture of this kernel (same N only)
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "ncu_r01_traffic.json")
        if os.path.exists(tpath):
            ent = json.load(open(tpath)).get(top)
            if ent and ent.get("npts") == v["units"] // max(1, v["launches"]):
                traffic = ent["dram_bytes_per_launch"]
        roofline = {"kernel": top, "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                    "traffic": traffic, "peak_source": peak_src, "avg_launch_ms": v["ms"] / v["launches"],
                    "algorithmic_bytes_per_launch": algorithmic_bytes(top, v["units"]) / v["launches"],
                    "points_per_launch": v["units"] / v["launches"]}
        fl = algorithmic_flops(top, v["units"])
        if fl is not None:
            # the spreading kernels are FP64-pipe bound, not HBM bound (ncu: FP64 pipe ~68 % busy, DRAM ~24 %):
            # report the FP64 fraction against a DFMA peak measured on this device in this run
            fp64_peak = _lib.measure_fp64_peak()
            roofline["fp64"] = {"achieved": fl / v["ms"] / 1e9, "peak": fp64_peak, "unit": "TFLOP/s",
                                "frac": fl / v["ms"] / 1e9 / fp64_peak, "flops_per_point": fl / v["units"],
                                "peak_source": "measured in this run (efgp_measure_fp64_peak: pure DFMA kernel)"}
        # the dominant kernel that IS bounded by HBM, for comparison
        hbm_own = {k: w for k, w in own.items() if algorithmic_flops(k, 1) is None}
        if hbm_own and top not in hbm_own:
            t2 = max(hbm_own, key=lambda k: hbm_own[k]["ms"])
            a2 = algorithmic_bytes(t2, hbm_own[t2]["units"]) / hbm_own[t2]["ms"] / 1e6
            roofline["dominant_hbm_bound_kernel"] = {"kernel": t2, "achieved": a2, "peak": peak, "unit": "GB/s",
                                                      "frac": a2 / peak,
                                                      "avg_launch_ms": hbm_own[t2]["ms"] / hbm_own[t2]["launches"]}
    launches = sum(v["launches"] for v in stages.values())

    line = {"metric": METRIC, "value": n_total / ms / 1e3, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, world, stats.get("mtot")),
            "numa_cpulist": numa, "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
            "stages": per_stage, "mean_cg_iters": stats.get("mean_cg_iters"), "result_rms": check}
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args.cpu_sample, 2)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
