"""Experiment: why do the SAME kernels run ~20 % slower in the cold step of bench.py than in the refit loop?
Variants differ only in host synchronisation / object lifetime."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gp-quadrature_b200")); sys.path.insert(0, ROOT)
import torch
import efgpnd
from efgp_b200 import _lib
from kernels.squared_exponential import SquaredExponential
import bench

dev = torch.device("cuda")
N = int(float(os.environ.get("EXP_N", "1e8")))
x, y = bench.synth_range(0, N, 2, dev, torch.float64)
ELL, VAR, SIG2, EPS = 0.05, 1.0, 0.1, 1e-4
keep = {}


def fresh(clear_plan, clear_bbox, sync, ell=ELL):
    if sync:
        torch.cuda.synchronize()
    if clear_plan:
        efgpnd._PLAN_CACHE.clear()
    if clear_bbox:
        efgpnd._BBOX_CACHE.clear()
    k = SquaredExponential(dimension=2, init_lengthscale=ell, init_variance=VAR)
    m = efgpnd.EFGPND(x, y, k, sigmasq=SIG2, eps=EPS, nufft_eps=1e-4, opts={"cg_tolerance": 1e-4}, estimate_params=False)
    return m.predict(x, return_variance=False)[0]


def persistent(sync, kstep):
    if sync:
        torch.cuda.synchronize()
    m = keep["m"]
    m.kernel.set_hyper("lengthscale", ELL * (1.0 + 0.02 * (1 + kstep % 2)))
    return m.predict(x, return_variance=False)[0]


def run(name, fn, steps=5):
    for k in range(3):
        fn(k)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(steps):
        fn(k)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    _lib.timing_enable(True)
    for k in range(steps):
        fn(k)
    torch.cuda.synchronize()
    st = _lib.timing_report()
    _lib.timing_enable(False)
    pick = {k: round(v["ms"] / steps, 3) for k, v in st.items() if v["ms"] / steps > 0.02}
    print(json.dumps({"variant": name, "ms_per_step": round(ms, 3), "stage_sum": round(sum(v["ms"] for v in st.values()) / steps, 3), **pick}), flush=True)


run("A cold: new plan + bbox sync", lambda k: fresh(True, True, False))
run("B new plan, bbox cached (no host sync)", lambda k: fresh(True, False, False))
run("B2 new plan, bbox cached, lengthscale 0.051", lambda k: fresh(True, False, False, 0.051))
run("B3 new plan, bbox cached, lengthscale 0.052", lambda k: fresh(True, False, False, 0.052))
run("B4 new plan, bbox cached, lengthscale 0.049", lambda k: fresh(True, False, False, 0.049))
k0 = SquaredExponential(dimension=2, init_lengthscale=ELL, init_variance=VAR)
keep["m"] = efgpnd.EFGPND(x, y, k0, sigmasq=SIG2, eps=EPS, nufft_eps=1e-4, opts={"cg_tolerance": 1e-4}, estimate_params=False)
keep["m"].predict(x, return_variance=False)
run("D refit persistent model (h changes)", lambda k: persistent(False, k))
def persistent_cold(kstep):
    keep["m"]._beta = None
    return persistent(False, kstep)


run("H refit persistent model, cold CG start", persistent_cold)
run("G new model each step at alternating lengthscale (plan kept, re-sorted), no sync",
    lambda k: fresh(False, False, False, ELL * (1.0 + 0.02 * (1 + k % 2))))
run("A2 cold again", lambda k: fresh(True, True, False))
