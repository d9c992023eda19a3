#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/q
mkdir -p $O
B="--steps 20 --warmup 5 --npts 1.25e7 --no-cpu-baseline --no-parity --no-e2e --no-f32"
run() { tag=$1; shift; env "$@" timeout 300 python bench.py $B > $O/bench_$tag.json 2> $O/bench_$tag.err; }
run auto A=1
run r3 EFGP_CG_CONV_R=3
run r4 EFGP_CG_CONV_R=4
timeout 600 python -m pytest tests/test_misc_gpu.py tests/test_model_gpu.py -q -m gpu 2>&1 | tail -1
python - <<'PY'
import json
for t in ["auto","r3","r4"]:
    try:
        d=json.load(open(f"gpurun_out/q/bench_{t}.json")); print(t, round(d["ms_per_step"],3), d["result_rms"], d["mean_cg_iters"], {k: round(v["ms_per_step"],4) for k,v in d["stages"].items() if "cg" in k})
    except Exception as e: print(t, "ERR", e)
PY
