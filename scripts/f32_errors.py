"""Print the actual float32 model errors against the reference's float32 golden (tests/test_model_gpu.py bars)."""
import contextlib, io, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import test_model_gpu as T
efgpnd, m, g, dt, neps = T.build("usa_f32")
x_new = torch.from_numpy(g["x_new"]).to(dt).cuda()
efgpnd.PROBE_DEVICE = "cpu"
with contextlib.redirect_stdout(io.StringIO()):
    mean, var = m.predict(x_new, return_variance=False)
    out = {"ws": T.rel(m._ws, g["ws"]), "beta": T.rel(m._beta, g["beta"]), "mean": T.rel(mean, g["mean"])}
    torch.manual_seed(321)
    _, var_s = m.predict(x_new, return_variance=True, variance_method="stochastic", hutchinson_probes=24)
    out["var_s"] = T.rel(var_s, g["var_stochastic"])
    _, var_r = m.predict(x_new[:32], return_variance=True, variance_method="regular")
    out["var_r"] = T.rel(var_r, g["var_regular"])
    efgpnd2, m2, g2, dt2, neps2 = T.build("usa_f32")
    torch.manual_seed(1234)
    grads = m2.compute_gradients(trace_samples=4, cg_tol=1e-5, nufft_eps=neps2, apply_gradients=True)
    out["grad"] = T.rel(grads, g2["grad_T4"])
    st = m2.last_gradient_stats
    out["iters"] = (int(st["mean_cg_iters"]), int(st["trace_cg_iters"]), [int(v) for v in g2["grad_T4_iters"]])
print(out)
