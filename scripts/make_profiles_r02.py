"""Build the round-2 profile artefacts under profiles/ from what a gpurun session brought back:

    python scripts/make_profiles_r02.py gpurun_out/final

expects in that directory: bench.json (1 GPU, default run), ncu_launches.csv (gpu__time_duration pass of the bench
command), ncu_full.ncu-rep (ncu --set full of the hot kernels); optional bench_shard.json, bench_2gpu.json,
bench_config4.json, bench_config4_2gpu.json, bench_reference.json.  Writes bench_r02_*.json, ncu_r02_launches.csv,
ncu_r02_summary.md, ncu_r02_traffic.json."""
import collections
import csv
import io
import json
import os
import re
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = os.path.join(ROOT, "profiles")
G = os.path.abspath(sys.argv[1]) if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "final")


def last_json(path):
    lines = [l for l in open(path).read().splitlines() if l.startswith("{")]
    return json.loads(lines[-1])


for src, dst in (("bench.json", "bench_r02_1gpu.json"), ("bench_shard.json", "bench_r02_shard.json"),
                 ("bench_2gpu.json", "bench_r02_2gpu.json"), ("bench_config4.json", "bench_r02_config4_1gpu.json"),
                 ("bench_config4_2gpu.json", "bench_r02_config4_2gpu.json"), ("bench_reference.json", "bench_r02_reference.json"),
                 ("bench_presort.json", "bench_r02_presort.json")):
    if os.path.exists(os.path.join(G, src)):
        json.dump(last_json(os.path.join(G, src)), open(os.path.join(P, dst), "w"), indent=1)

rep = os.path.join(G, "ncu_full.ncu-rep")
if not os.path.exists(rep):
    rep = os.path.join(G, "ncu_full_raw.csv")      # the raw page exported on the box (the report itself is > 64 MiB)
tab = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "ncu_summary.py"), rep], capture_output=True, text=True).stdout
raw = (open(rep).read() if rep.endswith(".csv") else
       subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout)
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
ki, ri, wi = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
stage_of = [("bucket_hist", "sort.bucket_hist_kernel"), ("bucket_scatter", "sort.bucket_scatter_kernel_pay"),
            ("bin_hist", "sort.bin_hist_kernel"), ("bin_scatter", "sort.bin_scatter_kernel_pay"),
            ("spread2d_moment_pair_kernel<9, 5", "spread2d_pair_w9_w5_f64"),
            ("spread2d_moment_kernel<9, 0", "spread2d_w9_f64_ones"), ("moment_to_grid_kernel", "type1.moment_to_grid"),
            ("spread2d_kernel<double, 9, 0", "spread2d_w9_f64_ones"), ("spread2d_kernel<double, 7, 0", "spread2d_w7_f64_ones"),
            ("spread2d_kernel<double, 5, 2", "spread2d_w5_f64_rec"), ("interp2d_kernel<double, 5", "interp2d_w5_f64"),
            ("slot_to_bucket", "unpermute.slot_to_bucket_order"), ("from_bucket", "unpermute.from_bucket_order"),
            ("bbox_partial", "fit.bbox"), ("cg_conv_kernel", "cg.iterations"), ("cg_persist_kernel", "cg.iterations")]
bench = last_json(os.path.join(G, "bench.json"))
npts = int(bench["config"]["N_total"])
traffic = {}
for r in rows[2:]:
    for pat, st in stage_of:
        if pat in r[ki] and st not in traffic:
            b = float(r[ri]) * scale[units[ri]] + float(r[wi]) * scale[units[wi]]
            traffic[st] = {"dram_bytes_per_launch": b, "npts": npts,
                           "source": "profiles/ncu_r02_summary.md (ncu --set full, one launch)"}
json.dump(traffic, open(os.path.join(P, "ncu_r02_traffic.json"), "w"), indent=1)

lines = open(os.path.join(G, "ncu_launches.csv")).read().splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
agg = collections.OrderedDict()
for r in csv.DictReader(lines[start:]):
    if r["Metric Name"] != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"]
    ms = v * {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}.get(u, 1e-6)
    name = re.split(r"\(", r["Kernel Name"])[0][:90]
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += ms
open(os.path.join(P, "ncu_r02_launches.csv"), "w").write("\n".join(lines[start:]) + "\n")
tot = sum(v[1] for v in agg.values())
ll = "| launches | total ms | share | kernel |\n|---|---|---|---|\n"
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:26]:
    ll += f"| {v[0]} | {v[1]:.3f} | {100 * v[1] / tot:.1f}% | `{k}` |\n"
st = bench["stages"]
ssum = sum(v["ms_per_step"] for v in st.values())
live = "| stage | live ms/step | share of stage sum | frac of HBM (algorithmic bytes) |\n|---|---|---|---|\n"
for k, v in sorted(st.items(), key=lambda kv: -kv[1]["ms_per_step"])[:14]:
    live += f"| {k} | {v['ms_per_step']:.3f} | {100 * v['ms_per_step'] / ssum:.1f}% | {v.get('hbm_frac', float('nan')):.2f} |\n"
md = f"""# Round 2 — ncu evidence (one B200, N = {npts:.0e}, d = 2, fp64)

Commands (B200_PROFILING.md recipe, run under gpurun after the same bench command exited 0 without ncu); regenerated by
`scripts/make_profiles_r02.py` from what the call brought back:

```
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:"efgp::|regular_fft|vector_fft|cub::" -c 900 --csv --log-file ncu_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-parity --no-e2e
ncu --set full --clock-control none --import-source on -k regex:"spread2d_|moment_to_grid|interp2d_kernel|bucket_|bin_|cg_conv|cg_persist|from_bucket|slot_to_bucket|bbox_partial" --launch-skip 36 -c 14 -o /tmp/ncu_full python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-parity --no-e2e
```

## Per-launch counters of one whole solve (`ncu --set full`, raw page)

{tab}

## Launch list (`gpu__time_duration.sum`, cold-cache and serialised: compare SHARES, not absolutes)

{ll}
total {tot:.2f} ms over the profiled launches (library, cuFFT and CUB kernels of the warm-up, timed and refit steps; the\nsynthetic-data ATen kernels are filtered out with -k)

## Live CUDA-event stage times of the same build (`profiles/bench_r02_1gpu.json`, `stages`)

step {bench['ms_per_step']:.3f} ms (uninstrumented) / {bench.get('ms_per_step_with_stage_events', float('nan')):.3f} ms (with stage events); stage sum {ssum:.3f} ms

{live}
"""
open(os.path.join(P, "ncu_r02_summary.md"), "w").write(md)
print(md[-2500:])
