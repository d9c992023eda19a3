#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/q
mkdir -p $O
timeout 2400 python -m pytest tests -q -m gpu --timeout=900 > $O/pytest.log 2>&1
grep -E "^FAILED|^E  |passed|failed" $O/pytest.log | head -20
B="--steps 5 --warmup 3 --no-cpu-baseline --no-parity --no-e2e --no-f32"
run() { tag=$1; shift; env "$@" timeout 300 python bench.py $B > $O/bench_$tag.json 2> $O/bench_$tag.err; }
run base A=1
B="--steps 20 --warmup 5 --npts 1.25e7 --no-cpu-baseline --no-parity --no-e2e --no-f32"
run shard A=1
python - <<'PY'
import json
for t in ["base","shard"]:
    try:
        d=json.load(open(f"gpurun_out/q/bench_{t}.json")); ss=sum(v["ms_per_step"] for v in d["stages"].values()); print(t, round(d["ms_per_step"],3), "stage sum", round(ss,3), "refit", round(d["refit"]["ms_per_step"],3), d["result_rms"])
    except Exception as e: print(t, "ERR", e)
PY
timeout 600 python scripts/timeline.py 1.25e7 8 2>&1 | tail -22 | head -12
