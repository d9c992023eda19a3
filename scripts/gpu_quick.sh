#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out/q
timeout 600 python -m pytest tests/test_misc_gpu.py tests/test_model_gpu.py -q -m gpu --timeout=600 > gpurun_out/q/pytest_conv.log 2>&1
grep -E "^FAILED|^E  |passed|failed" gpurun_out/q/pytest_conv.log | head -30
B="--steps 5 --warmup 3 --no-cpu-baseline --no-parity --no-e2e"
timeout 300 python bench.py $B > gpurun_out/q/bench_conv.json 2> gpurun_out/q/bench_conv.err
EFGP_CG_CONV=0 timeout 300 python bench.py $B > gpurun_out/q/bench_noconv.json 2> gpurun_out/q/bench_noconv.err
timeout 300 python bench.py --steps 20 --warmup 5 --npts 1.25e7 --no-cpu-baseline --no-parity --no-e2e > gpurun_out/q/bench_conv_shard.json 2> gpurun_out/q/bench_conv_shard.err
python - <<'PY'
import json
for t in ["conv","noconv","conv_shard"]:
    try:
        d=json.load(open(f"gpurun_out/q/bench_{t}.json")); print(t, d["ms_per_step"], d["stages"]["cg.iterations"], d["mean_cg_iters"], d["result_rms"])
    except Exception as e: print(t, "ERR", e)
PY
