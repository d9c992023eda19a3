#!/bin/bash
# Round-2 GPU session 2 (1 GPU): full parity suite, headline bench (uninstrumented value), config-4 bench, presorted-input run.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/s3
mkdir -p $O
echo "== pytest" ; date
timeout 2400 python -m pytest tests -q -m gpu --timeout=1200 --deselect tests/test_nccl_sharding_gpu.py > $O/pytest.log 2>&1
echo "pytest rc=$?" | tee -a $O/pytest.log
tail -15 $O/pytest.log
echo "== bench" ; date
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err
echo "bench rc=$?"; tail -c 800 $O/bench.err
echo "== bench sigma5" ; date
EFGP_SIGMA=5 timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-parity --no-e2e > $O/bench_sigma5.json 2> $O/bench_sigma5.err
echo "== bench presorted" ; date
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-parity --no-e2e --presort > $O/bench_presort.json 2> $O/bench_presort.err
echo "== bench N/8 (the 8-GPU shard size on one GPU)" ; date
timeout 300 python bench.py --steps 20 --warmup 5 --npts 1.25e7 --no-cpu-baseline --no-parity --no-e2e > $O/bench_1p25e7.json 2> $O/bench_1p25e7.err
echo "== config4" ; date
timeout 900 python bench.py --workload config4 --steps 2 --warmup 3 > $O/bench_config4.json 2> $O/bench_config4.err
echo "config4 rc=$?"; tail -c 800 $O/bench_config4.err
ls -la $O
date
