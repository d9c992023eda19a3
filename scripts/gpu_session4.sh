#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/s4
mkdir -p $O
echo "== pytest" ; date
timeout 2400 python -m pytest tests -q -m gpu --timeout=1200 --deselect tests/test_nccl_sharding_gpu.py > $O/pytest.log 2>&1
echo "pytest rc=$?" | tee -a $O/pytest.log
grep -E "^FAILED|^E  |passed|failed" $O/pytest.log | head -40
echo "== exp_sync" ; date
timeout 600 python scripts/exp_sync.py > $O/exp_sync.log 2>&1
cat $O/exp_sync.log | grep variant
echo "== bench" ; date
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err
echo "bench rc=$?"; tail -c 600 $O/bench.err
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-parity --no-e2e --presort > $O/bench_presort.json 2> $O/bench_presort.err
timeout 900 python bench.py --workload config4 --steps 2 --warmup 3 > $O/bench_config4.json 2> $O/bench_config4.err
date
