#!/bin/bash
# no-alloc entry point alone in a fresh process, then the whole GPU suite (it runs last), then the conv-CG A/B
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/s7
mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu --timeout=600 > $O/pytest_noalloc_alone.log 2>&1
echo "alone rc=$?"; tail -3 $O/pytest_noalloc_alone.log
timeout 2400 python -m pytest tests -q -m gpu --timeout=900 > $O/pytest.log 2>&1
echo "suite rc=$?"; grep -E "^FAILED|^E  |passed|failed" $O/pytest.log | head -30
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; tail -1 $O/smoke.log
B="--steps 5 --warmup 3 --no-cpu-baseline --no-parity --no-e2e"
timeout 300 python bench.py $B > $O/bench_conv.json 2> $O/bench_conv.err
EFGP_CG_CONV=0 timeout 300 python bench.py $B > $O/bench_noconv.json 2> $O/bench_noconv.err
EFGP_SPREAD_MOMENTS=0 timeout 300 python bench.py $B > $O/bench_nomom.json 2> $O/bench_nomom.err
EFGP_SPREAD_MOMENTS=5 timeout 300 python bench.py $B > $O/bench_mom5.json 2> $O/bench_mom5.err
timeout 300 python bench.py --steps 20 --warmup 5 --npts 1.25e7 --no-cpu-baseline --no-parity --no-e2e > $O/bench_conv_shard.json 2> $O/bench_conv_shard.err
python - <<'PY'
import json
for t in ["conv","noconv","nomom","mom5","conv_shard"]:
    try:
        d=json.load(open(f"gpurun_out/s7/bench_{t}.json")); print(t, d["ms_per_step"], d["stages"]["cg.iterations"], d["mean_cg_iters"], d["result_rms"], {k: round(v["ms_per_step"],3) for k,v in d["stages"].items() if "spread" in k})
    except Exception as e: print(t, "ERR", e)
PY
