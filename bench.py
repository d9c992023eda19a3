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

`--impl reference` times the reference's CPU path on a bounded sample of the same workload, on all host
cores: the UNMODIFIED reference modules (oracle/_ref, copied there by oracle/make_ref.sh) with the C/OpenMP
ES-kernel NUFFT of oracle/cnufft.c under the `pytorch_finufft` name (FINUFFT is not installable here); if
oracle/_ref is missing it falls back to the oracle's restatement of efgpnd.py (kind "port").

`--workload config4` times BASELINE.json configs[3] instead: one 3-D SE hyperparameter-gradient step at N = 1e7
with 64 Hutchinson probes, the probes sharded over the GPUs (scalar all-reduce).
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
    ap.add_argument("--cpu-sample", type=float, default=1e7, help="points in the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-f32", action="store_true", help="skip the extra float32 figure")
    ap.add_argument("--no-parity", action="store_true", help="skip the full-size oracle comparison after the timing")
    ap.add_argument("--workload", default="config3", choices=["config3", "config4"])
    ap.add_argument("--probes", type=int, default=64, help="config4: Hutchinson probes per gradient step")
    ap.add_argument("--presort", action="store_true",
                    help="spatially presort each rank's points once before the timing (utils.datasets.spatial_presort: the "
                         "ingestion step of SURVEY 8f-3); the timed step is unchanged and still sorts by fine-grid cell")
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


CHUNK = 1 << 20


def synth_range(lo, hi, d, device, dtype, seed=1000):
    """Points [lo, hi) of ONE seeded data set (chunk c of 2^20 points comes from seed + c), so that every rank of
    an N-GPU run holds a slice of the data the 1-GPU run solves and the results can be compared."""
    import torch
    xs, ys = [], []
    for c in range(lo // CHUNK, (max(hi, lo + 1) - 1) // CHUNK + 1):
        x, y = synth(CHUNK, d, seed + c, device, dtype)
        a, b = max(lo, c * CHUNK) - c * CHUNK, min(hi, (c + 1) * CHUNK) - c * CHUNK
        xs.append(x[a:b])
        ys.append(y[a:b])
    return torch.cat(xs), torch.cat(ys)


# ----------------------------------------------------------------------------- CPU reference arm
def _host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


REF_DIR = os.path.join(ROOT, "oracle", "_ref")


def _cpu_arm_setup():
    """Returns (kind, solve) for the CPU arm.  kind "reference": the UNMODIFIED reference modules in oracle/_ref
    (EFGPND(...).predict, efgpnd.py:824-972) with the C/OpenMP NUFFT under the pytorch_finufft name; kind "port": the
    oracle's restatement of the same algebra when oracle/_ref is absent.  Must run in a process that has NOT imported
    the product package (same module names)."""
    import torch
    threads = _host_threads()
    torch.set_num_threads(threads)            # also under torchrun, which exports OMP_NUM_THREADS=1
    oracle_dir = os.path.join(ROOT, "oracle")
    if os.path.exists(os.path.join(REF_DIR, "efgpnd.py")):
        for mod in ("efgpnd", "cg", "kernels", "utils"):
            assert mod not in sys.modules, f"{mod} already imported: the reference arm needs its own process"
        sys.path[:] = [pth for pth in sys.path if os.path.abspath(pth or ".") not in (PKG, ROOT)]
        sys.path.insert(0, REF_DIR)
        sys.path.insert(0, os.path.join(oracle_dir, "fast_shim"))
        sys.path.append(oracle_dir)
        import fast_nufft as F
        F.NTHREADS = threads
        import io
        import contextlib
        import warnings
        warnings.filterwarnings("ignore")
        import efgpnd as R
        from kernels.squared_exponential import SquaredExponential

        def solve(x, y):
            kern = SquaredExponential(dimension=2, init_lengthscale=ELL, init_variance=VAR)
            with contextlib.redirect_stdout(io.StringIO()):
                m = R.EFGPND(x, y, kern, sigmasq=SIG2, eps=EPS, nufft_eps=NUFFT_EPS, opts={"cg_tolerance": CG_TOL},
                             estimate_params=False)
                mean, _ = m.predict(x, return_variance=False)
            return int(round(m._xis.shape[0] ** 0.5)), mean

        return "reference", solve
    sys.path.insert(0, oracle_dir)
    import efgp_oracle as O
    import fast_nufft as F
    F.NTHREADS = threads

    def solve(x, y):
        kern = O.OracleKernel("se", 2, ELL, VAR)
        f = O.fit(x, y, kern, SIG2, EPS, NUFFT_EPS, cg_tol=CG_TOL, backend=F)
        return f.mtot, O.predict_mean(f, x, NUFFT_EPS, backend=F)

    return "port", solve


def run_reference(args):
    """CPU arm: K timed posterior-mean solves of a bounded sample of the workload on all host cores, preceded by a
    size sweep (1e6 / 4e6 / sample) that shows the metric (Mpts/s) is size-independent for N >> M, which is what lets a
    sample stand in for N = 1e8."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    kind, solve = _cpu_arm_setup()
    reps = max(1, args.steps)
    n_s = int(args.cpu_sample)
    x, y = synth(n_s, 2, 1234, "cpu", torch.float64)
    for _ in range(max(1, min(args.warmup, 2))):
        mtot, _ = solve(x[:200000], y[:200000])
    sweep = {}
    for n in sorted({min(n_s, 1000000), min(n_s, 4000000)}):
        if n < n_s:
            t0 = time.perf_counter()
            solve(x[:n], y[:n])
            sweep[str(n)] = n / (time.perf_counter() - t0) / 1e6
    t0 = time.perf_counter()
    for _ in range(reps):
        mtot, _ = solve(x, y)
    dt = (time.perf_counter() - t0) / reps
    v = n_s / dt / 1e6
    sweep[str(n_s)] = v
    cores = _host_threads()
    what = ("UNMODIFIED reference EFGPND.predict (oracle/_ref) + C/OpenMP ES-kernel NUFFT as pytorch_finufft"
            if kind == "reference" else "oracle restatement of efgpnd.py + C/OpenMP ES-kernel NUFFT")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": reps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, 1, mtot),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": f"{n_s} points per step (bounded sample of the N={int(args.npts)} workload); "
                                       f"{what}; FINUFFT itself is not installable here",
                             "mpts_per_s_by_sample_size": sweep, "os_cpu_count": os.cpu_count()},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def cpu_baseline(args):
    """The CPU arm in its own process (the reference's modules share their names with the product package)."""
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "2", "--warmup", "1",
           "--cpu-sample", str(args.cpu_sample), "--npts", str(args.npts)]
    env = dict(os.environ)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "OMP_NUM_THREADS"):
        env.pop(k, None)
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=900)
    for ln in reversed(r.stdout.splitlines()):
        if ln.startswith("{"):
            d = json.loads(ln)
            cb = d["cpu_baseline"]
            cb["seconds_per_step"] = d["ms_per_step"] / 1e3
            return cb
    return {"value": None, "unit": UNIT, "cores": _host_threads(), "kind": "failed",
            "sample": (r.stderr or r.stdout)[-400:]}


def workload_config(args, world, mtot=None):
    return {"workload": "BASELINE.json configs[2]: 2-D squared-exponential EFGP posterior-mean solve "
                        "(sort + F^*1 @6e-8 + F^*y + Toeplitz + PCG + mean at the N training points)",
            "N_total": int(args.npts), "d": 2, "lengthscale": ELL, "variance": VAR, "sigmasq": SIG2, "eps": EPS,
            "nufft_eps": NUFFT_EPS, "cg_tolerance": CG_TOL, "mtot": mtot,
            "parallelism": f"points sharded over {world} GPU(s), NCCL all-reduce of the mode grids" if world > 1
            else "single GPU",
            "point_order": "spatially presorted (Morton order, once, untimed)" if getattr(args, "presort", False)
            else "random (as generated)",
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
        self._sample()                    # probe that the queries work; not part of the statistics
        self.sm.clear()
        self.pw.clear()
        self.reasons.clear()
        self.t = threading.Thread(target=self._run, daemon=True)

    def start(self):
        """Called right before the first timed launch: every sample lies inside the timed region."""
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
        sm = sorted(self.sm)                      # every sample taken between start and stop (the timed region)
        return {"sm_mhz": sm[len(sm) // 2], "sm_mhz_min": sm[0], "sm_max_mhz": self.mx, "reasons": sorted(self.reasons),
                "samples": len(sm), "power_w_max": max(self.pw) if self.pw else None, "source": "nvml"}


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
        self.gpu_index = gpu_index

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={self.Q}",
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
        sm = sorted(sm)                           # every sample of the sampling window
        return {"sm_mhz": sm[len(sm) // 2], "sm_mhz_min": sm[0], "sm_max_mhz": max(mx), "reasons": sorted(reasons),
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
    """FP64 flops (FMA = 2) the spreading / interpolation kernels need for n points.  Register-footprint kernels
    (spread.cu es_eval + W^d FMAs): per dimension the even/odd Horner evaluation of the W piecewise polynomials of
    degree W - 1 (W for W <= 4: common.cuh es_degree), then W^d FMAs.  The 2-D fp64 spreads of width >= 7 on dense
    point sets (every bench workload) run the moment kernel (spread2d_moment_kernel): 2 (NC - 1) multiplications for the
    powers of the two in-cell offsets and NC^2 FMAs per point, NC = degree + 1."""
    import re
    mp = re.match(r"spread2d_pair_w(\d+)_w(\d+)_f64", stage)
    if mp:      # fused generator + right-hand side (spread2d_moment_pair_kernel): powers once, two moment sets
        ng, nr = int(mp.group(1)), int(mp.group(2))          # NC = W for the odd widths >= 5 (degree W - 1)
        return n * (2 * (ng - 1) + nr + 2 * (ng * ng + nr * nr) + 4)
    m = re.match(r"(spread|interp)(\d)d_w(\d+)_(f64|f32)(.*)", stage)
    if not m:
        return None
    kind, W, prec, rest = m.group(1), int(m.group(3)), m.group(4), m.group(5)
    deg = W if W <= 4 else W - 1
    if kind == "spread" and d == 2 and prec == "f64" and W >= 7 and "probes" not in rest:
        nc = deg + 1
        return n * (2 * (nc - 1) + 2 * nc * nc + 4 + (0 if rest == "_ones" else nc))
    ne, no = deg // 2 + 1, (deg + 1) // 2
    horner = (W // 2) * (2 * (ne + no - 2) + 3) + (2 * (ne - 1) if W % 2 else 0) + 4
    return n * (d * horner + 2 * W ** d + W)


def parity_check(x_loc, beta_timed, mean_timed, beta_tight, mean_tight, n_total, world, dev):
    """Rank 0, after the timing: the CPU oracle (oracle/efgp_oracle.py restating efgpnd.py:745-822, on the C/OpenMP
    ES-kernel NUFFT) solves the SAME full-size problem; beta is compared directly and the posterior mean at 10^4
    sampled points against the exact NUDFT of the oracle's beta.  Two configurations: the timed one (nufft_eps 1e-4 on
    both sides: bar 10 * nufft_eps, BASELINE.json) and a tight one (nufft_eps 1e-9, CG 1e-10: bar 1e-6 on the posterior
    mean, BASELINE.json's tolerance)."""
    import math
    import torch
    odir = os.path.join(ROOT, "oracle")
    if odir not in sys.path:
        sys.path.append(odir)
    import efgp_oracle as O
    import fast_nufft as F
    from pytorch_finufft import functional as exact
    F.NTHREADS = _host_threads()
    torch.set_num_threads(_host_threads())
    t_start = time.perf_counter()
    if world > 1:
        xa, ya = synth_range(0, n_total, 2, dev, torch.float64)
        x_all, y_all = xa.cpu(), ya.cpu()
        del xa, ya
    else:
        x_all, y_all = x_loc.cpu(), None
    if y_all is None:
        _, ya = synth_range(0, n_total, 2, dev, torch.float64)
        y_all = ya.cpu()
        del ya
    kern = O.OracleKernel("se", 2, ELL, VAR, param_dtype=torch.float64)
    s2 = float(torch.exp(torch.log(torch.tensor(SIG2, dtype=torch.float64))))
    g = torch.Generator().manual_seed(7)
    idx = torch.randint(0, x_loc.shape[0], (10000,), generator=g)
    xs = x_loc[idx.to(dev)].cpu()

    def rel(a, b):
        return float(torch.linalg.norm((a - b).reshape(-1)) / torch.linalg.norm(b.reshape(-1)))

    out = {"n_checked_points": int(n_total), "n_checked_means": int(idx.numel()),
           "oracle": "oracle/efgp_oracle.fit on oracle/cnufft.c (CPU, all host cores) at the full N; means against the "
                     "exact NUDFT of the oracle's beta at sampled points"}
    for tag, neps, ctol, beta, mean, bar in (("timed_config", NUFFT_EPS, CG_TOL, beta_timed, mean_timed, 10 * NUFFT_EPS),
                                             ("tight", 1e-9, 1e-10, beta_tight, mean_tight, 1e-6)):
        f = O.fit(x_all, y_all, kern, s2, EPS, neps, cg_tol=ctol, backend=F)
        phi = (2 * math.pi * f.h * xs).T.contiguous()
        mean_ref = exact.finufft_type2(phi, (f.ws * f.beta).reshape(f.mtot, f.mtot), isign=+1, modeord=False).real
        bc = beta.cpu().reshape(-1).to(f.rhs.dtype)
        # beta itself is only determined up to (CG tolerance) x (condition number) in the directions the weights
        # damp; what the reference's solve guarantees is the RESIDUAL of its system, so our beta is put into the
        # ORACLE's operator and right-hand side: ||A_oracle beta_ours - rhs_oracle|| / ||rhs_oracle||
        out[tag] = {"nufft_eps": neps, "cg_tol": ctol,
                    "system_residual_rel": rel(f.A_mean(bc), f.rhs),
                    "oracle_own_residual_rel": rel(f.A_mean(f.beta.reshape(-1)), f.rhs),
                    "beta_rel": rel(bc, f.beta.reshape(-1)),
                    "weighted_beta_rel": rel(f.ws.reshape(-1) * bc, (f.ws * f.beta).reshape(-1)),
                    "mean_rel": rel(mean[idx.to(dev)].cpu(), mean_ref), "bar": bar, "oracle_cg_iters": int(f.iters)}
        out[tag]["ok"] = bool(out[tag]["mean_rel"] <= bar and out[tag]["system_residual_rel"] <= max(bar, 10 * ctol))
    out["seconds"] = time.perf_counter() - t_start
    return out


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


# ----------------------------------------------------------------------------- config 4: probe-sharded gradient step
C4 = {"N": 1e7, "d": 3, "ell": 0.1, "var": 1.0, "sig2": 0.04, "eps": 1e-4, "cg_tol": 1e-5, "nufft_eps": 1e-5}
C4_METRIC, C4_UNIT = "efgp_gradient_step_probe_throughput", "Mprobe-pts/s"


def c4_data(n, device, dtype):
    import torch
    g = torch.Generator(device=device).manual_seed(0)                 # the same points on every rank
    x = torch.rand(n, 3, dtype=dtype, device=device, generator=g) * 2 - 1
    y = (torch.sin(3 * x[:, 0]) * torch.cos(4 * x[:, 1]) + 0.3 * torch.sin(5 * x[:, 2])
         + 0.2 * torch.randn(n, dtype=dtype, device=device, generator=g))
    return x, y


def c4_config(args, n, world):
    return {"workload": "BASELINE.json configs[3]: 3-D squared-exponential hyperparameter-gradient step "
                        "(EFGPND.compute_gradients: mean solve + Hutchinson trace with batched NUFFTs and batched PCG)",
            "N_total": int(n), "d": 3, "probes": int(args.probes), "lengthscale": C4["ell"], "variance": C4["var"],
            "sigmasq": C4["sig2"], "eps": C4["eps"], "nufft_eps": C4["nufft_eps"], "cg_tolerance": C4["cg_tol"],
            "parallelism": f"probes sharded over {world} GPU(s), every rank holds all points, scalar all-reduce"
            if world > 1 else "single GPU",
            "l2_policy": "inputs (320 MB of points, 5.8M-cell fine grids per probe) exceed the 126 MB L2"}


def run_config4_reference(args):
    """CPU arm of config 4: the reference's compute_gradients on a bounded sample (N = 2e5 points, 4 probes)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import torch
    kind, _ = _cpu_arm_setup()
    n_s, T = 200000, 4
    x, y = c4_data(n_s, "cpu", torch.float64)
    if kind == "reference":
        import io
        import contextlib
        import efgpnd as R
        from kernels.squared_exponential import SquaredExponential

        def step():
            k = SquaredExponential(dimension=3, init_lengthscale=C4["ell"], init_variance=C4["var"])
            with contextlib.redirect_stdout(io.StringIO()):
                m = R.EFGPND(x, y, k, sigmasq=C4["sig2"], eps=C4["eps"], estimate_params=False)
                torch.manual_seed(1234)
                m.compute_gradients(trace_samples=T, cg_tol=C4["cg_tol"], nufft_eps=C4["nufft_eps"], apply_gradients=False)
    else:
        import efgp_oracle as O
        import fast_nufft as F

        def step():
            k = O.OracleKernel("se", 3, C4["ell"], C4["var"])
            torch.manual_seed(1234)
            O.gradient(x, y, k, C4["sig2"], C4["eps"], T, C4["nufft_eps"], C4["cg_tol"], backend=F)
    step()
    reps = max(1, min(args.steps, 3))
    t0 = time.perf_counter()
    for _ in range(reps):
        step()
    dt = (time.perf_counter() - t0) / reps
    v = n_s * T / dt / 1e6
    print(json.dumps({"impl": "reference", "metric": C4_METRIC, "value": v, "unit": C4_UNIT, "n_gpus": args.gpus,
                      "steps": reps, "warmup": 1, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong",
                      "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": c4_config(args, C4["N"], 1),
                      "cpu_baseline": {"value": v, "unit": C4_UNIT, "cores": _host_threads(), "kind": kind,
                                       "sample": f"{n_s} points x {T} probes per step (bounded sample of N=1e7 x 64)"},
                      "e2e": {"value": v, "unit": C4_UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                      "gpu_launches": 0}), flush=True)


def run_config4(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback for the product path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import efgpnd
    from efgp_b200 import _lib
    from kernels.squared_exponential import SquaredExponential
    n = int(C4["N"]) if args.npts == 1e8 else int(args.npts)
    T = int(args.probes)
    x, y = c4_data(n, dev, torch.float64)
    opts = {"shard_probes": world > 1}
    state = {}

    def step(xd, yd):
        k = SquaredExponential(dimension=3, init_lengthscale=C4["ell"], init_variance=C4["var"])
        m = efgpnd.EFGPND(xd, yd, k, sigmasq=C4["sig2"], eps=C4["eps"], estimate_params=False, opts=opts)
        torch.manual_seed(1234)                       # identical probe seeds on every rank and repeat
        g = m.compute_gradients(trace_samples=T, cg_tol=C4["cg_tol"], nufft_eps=C4["nufft_eps"], apply_gradients=False)
        state["stats"] = m.last_gradient_stats
        return g

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        grads = step(x, y)
    barrier()
    sampler = make_clock_sampler(local) if rank == 0 else None
    _lib.timing_enable(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.reset_peak_memory_stats()
    barrier()
    if sampler:
        sampler.start()
    e0.record()
    for _ in range(args.steps):
        grads = step(x, y)
    e1.record()
    barrier()
    clocks = sampler.stop() if sampler else None
    stages = _lib.timing_report()
    _lib.timing_enable(False)
    tt = torch.tensor([e0.elapsed_time(e1) / args.steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms = float(tt.item())
    # end to end: points and targets come from pinned host memory every step, the gradient goes back to the host
    xh = torch.empty(x.shape, dtype=x.dtype, pin_memory=True).copy_(x)
    yh = torch.empty(y.shape, dtype=y.dtype, pin_memory=True).copy_(y)
    gh = torch.empty(3, dtype=torch.float64, pin_memory=True)
    xd, yd = torch.empty_like(x), torch.empty_like(y)

    def e2e_step():
        xd.copy_(xh, non_blocking=True)
        yd.copy_(yh, non_blocking=True)
        gh.copy_(step(xd, yd).to(torch.float64), non_blocking=True)

    e2e_step()
    barrier()
    e0.record()
    for _ in range(args.steps):
        e2e_step()
    e1.record()
    barrier()
    tt = torch.tensor([e0.elapsed_time(e1) / args.steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms_e2e = float(tt.item())
    if rank == 0:
        st = state["stats"]
        per_stage = {k: {"ms_per_step": v["ms"] / args.steps, "launches_per_step": v["launches"] / args.steps}
                     for k, v in stages.items()}
        n_side = sum(v["ms"] for k, v in stages.items() if k.split(".")[0] in ("sort", "permute", "unpermute", "type1", "type2")
                     or k.startswith("spread") or k.startswith("interp")) / args.steps
        cg_side = sum(v["ms"] for k, v in stages.items() if k.startswith("cg.") or k.startswith("toeplitz.")) / args.steps
        line = {"metric": C4_METRIC, "value": n * T / ms / 1e3, "unit": C4_UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": c4_config(args, n, world),
                "clocks": clocks,
                "e2e": {"value": n * T / ms_e2e / 1e3, "unit": C4_UNIT, "ms_per_step": ms_e2e,
                        "h2d_bytes_per_step": (xh.numel() + yh.numel()) * 8 * world, "d2h_bytes_per_step": 24 * world,
                        "api": "EFGPND(x, y, SquaredExponential).compute_gradients(trace_samples=64) on buffers copied "
                               "from pinned host memory; the 3 gradient entries copied back"},
                "gpu_launches": int(sum(v["launches"] for v in stages.values())),
                "split_ms_per_step": {"O(N) stages (sort, permutes, spread, interp, fine-grid FFTs)": n_side,
                                      "M-grid stages (Toeplitz, batched CG)": cg_side},
                "stages": per_stage, "mean_cg_iters": int(st["mean_cg_iters"]), "trace_cg_iters": int(st["trace_cg_iters"]),
                "mtot": int(st["mtot"]), "grad": [float(v) for v in grads],
                "peak_mem_GB": torch.cuda.max_memory_allocated() / 1e9}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        return run_config4_reference(args) if args.workload == "config4" else run_reference(args)
    if args.workload == "config4":
        return run_config4(args)
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
    x, y = synth_range(lo_pt, hi_pt, 2, dev, torch.float64)      # this rank's slice of ONE seeded data set
    if args.presort:
        from utils.datasets import spatial_presort
        x, y, _ = spatial_presort(x, y)
        torch.cuda.empty_cache()
    opts = {"cg_tolerance": CG_TOL, "shard_points": world > 1}
    stats = {}
    last = {}

    def solve(xd, yd, nufft_eps=NUFFT_EPS, cg_tol=CG_TOL):
        efgpnd._PLAN_CACHE.clear()                 # a cold step starts from raw points: bounding box and sort included
        efgpnd._BBOX_CACHE.clear()
        kern = SquaredExponential(dimension=2, init_lengthscale=ELL, init_variance=VAR)
        model = efgpnd.EFGPND(xd, yd, kern, sigmasq=SIG2, eps=EPS, nufft_eps=nufft_eps,
                              opts=dict(opts, cg_tolerance=cg_tol), estimate_params=False)
        mean, _ = model.predict(xd, return_variance=False)
        stats.update(model.last_fit_stats)
        last["model"] = model
        return mean

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        mean = solve(x, y)
    barrier()

    # ---- device-resident timing (value): K steps between two events, nothing else on the stream
    sampler = make_clock_sampler(local) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    if sampler:
        sampler.start()
    e0.record()
    for _ in range(args.steps):
        mean = solve(x, y)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1) / args.steps
    clocks = sampler.stop() if sampler else None      # stopped right after the closing synchronize
    # ---- the same K steps again with the library's per-stage CUDA events switched on (kernel durations for the
    # roofline and the launch count; the event pairs cost a little, which is why `value` is timed without them)
    _lib.timing_enable(True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        mean = solve(x, y)
    e1.record()
    barrier()
    ms_instrumented = e0.elapsed_time(e1) / args.steps
    stages = _lib.timing_report()
    _lib.timing_enable(False)
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    check = float(mean.double().square().mean().sqrt().item())

    # ---- refit after a hyperparameter change on the SAME data (what every optimiser step of the reference's loop
    # does, efgpnd.py:1138-1195): the model object, its library handle and everything that does not depend on the
    # hyperparameters (the bounding box; the sort when the grid spacing is unchanged) are kept
    kern_r = SquaredExponential(dimension=2, init_lengthscale=ELL, init_variance=VAR)
    model_r = efgpnd.EFGPND(x, y, kern_r, sigmasq=SIG2, eps=EPS, nufft_eps=NUFFT_EPS, opts=opts, estimate_params=False)
    model_r.predict(x, return_variance=False)

    def refit(k):
        model_r.kernel.set_hyper("lengthscale", ELL * (1.0 + 0.02 * (1 + k % 2)))
        return model_r.predict(x, return_variance=False)[0]

    for k in range(3):
        refit(k)
    barrier()
    e0.record()
    for k in range(args.steps):
        refit(3 + k)               # (continues the alternation: every timed step really changes the hypers)
    e1.record()
    barrier()
    tr = torch.tensor([e0.elapsed_time(e1) / args.steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tr, op=dist.ReduceOp.MAX)
    _lib.timing_enable(True)
    barrier()
    for k in range(args.steps):
        refit(3 + args.steps + k)
    barrier()
    refit_stages = {k: round(v["ms"] / args.steps, 4) for k, v in _lib.timing_report().items()}
    _lib.timing_enable(False)
    refit_info = {"ms_per_step": float(tr.item()), "value": n_total / float(tr.item()) / 1e3, "unit": UNIT,
                  "stages_ms": refit_stages, "mean_cg_iters": int(model_r.last_fit_stats["mean_cg_iters"]),
                  "mtot": int(model_r.last_fit_stats["mtot"]),
                  "what": "predict() after a lengthscale change on the same EFGPND object (same points): no bounding-box "
                          "read-back; the spacing h changes with the lengthscale, so the cell sort is redone"}
    del model_r

    # ---- the same cold step in float32 (an extra figure, not the headline: BASELINE.json's metric is fp64).  Inputs,
    # records, grids and CG vectors are float32 (half the sort / record traffic); the result is compared with the fp64
    # posterior mean of the timed configuration on the same points.
    f32_info = None
    if world == 1 and not args.no_f32:
        x32, y32 = x.float(), y.float()
        for _ in range(3):
            m32 = solve(x32, y32)
        barrier()
        e0.record()
        for _ in range(args.steps):
            m32 = solve(x32, y32)
        e1.record()
        barrier()
        ms32 = e0.elapsed_time(e1) / args.steps
        _lib.timing_enable(True)
        barrier()
        for _ in range(args.steps):
            m32 = solve(x32, y32)
        barrier()
        st32 = {k: round(v["ms"] / args.steps, 4) for k, v in _lib.timing_report().items()}
        _lib.timing_enable(False)
        rel32 = float((torch.linalg.norm(m32.double() - mean.double()) / torch.linalg.norm(mean.double())).item())
        f32_info = {"ms_per_step": ms32, "value": n_total / ms32 / 1e3, "unit": UNIT, "dtype": "f32",
                    "mean_rel_vs_f64": rel32, "mean_cg_iters": stats.get("mean_cg_iters"), "stages_ms": st32,
                    "what": "same cold step with float32 points, targets and arithmetic (nufft_eps, cg tolerance as "
                            "timed); mean_rel_vs_f64 = |mean32 - mean64| / |mean64| over all N points"}
        del x32, y32, m32
        mean = solve(x, y)             # leave the fp64 state (stats, last model) behind for what follows

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

    # ---- parity at the full size: every rank runs the two solves (collectives), rank 0 compares with the oracle
    parity = None
    if not args.no_parity:
        mean_timed = solve(x, y)
        beta_timed = last["model"]._beta.detach().clone()
        mean_tight = solve(x, y, nufft_eps=1e-9, cg_tol=1e-10)
        beta_tight = last["model"]._beta.detach().clone()
        barrier()
        if rank == 0:
            parity = parity_check(x, beta_timed, mean_timed, beta_tight, mean_tight, n_total, world, dev)
    beta_now = last["model"]._beta.detach()
    gprobe = torch.Generator(device="cpu").manual_seed(99)
    wprobe = torch.randn(beta_now.numel(), dtype=torch.float64, generator=gprobe).to(dev)
    beta_fingerprint = {"l2": float(beta_now.abs().square().sum().sqrt().item()),
                        "probe": float((beta_now.real * wprobe).sum().item())}

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
        tpath = os.path.join(ROOT, "profiles", "ncu_r02_traffic.json")
        if not os.path.exists(tpath):
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
            "stages": per_stage, "mean_cg_iters": stats.get("mean_cg_iters"), "result_rms": check,
            "beta_fingerprint": beta_fingerprint, "parity": parity, "refit": refit_info, "f32": f32_info,
            "ms_per_step_with_stage_events": ms_instrumented}
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
