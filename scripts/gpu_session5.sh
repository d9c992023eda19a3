#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/s5
mkdir -p $O
echo "== pytest" ; date
timeout 2400 python -m pytest tests -q -m gpu --timeout=1200 --deselect tests/test_nccl_sharding_gpu.py > $O/pytest.log 2>&1
echo "pytest rc=$?" | tee -a $O/pytest.log
grep -E "^FAILED|^E  |passed|failed" $O/pytest.log | head -30
B="--steps 5 --warmup 3 --no-cpu-baseline --no-parity --no-e2e"
echo "== bench default"; date
timeout 600 python bench.py $B > $O/bench_default.json 2> $O/bench_default.err
echo "== TMA interp"; date
EFGP_INTERP_TMA=1 timeout 300 python bench.py $B > $O/bench_tma.json 2> $O/bench_tma.err; echo "rc=$?"
echo "== max P 128"; date
EFGP_MAX_P=128 timeout 300 python bench.py $B > $O/bench_p128.json 2> $O/bench_p128.err
EFGP_MAX_P=512 timeout 300 python bench.py $B > $O/bench_p512.json 2> $O/bench_p512.err
echo "== N/8"; date
timeout 300 python bench.py --steps 20 --warmup 5 --npts 1.25e7 --no-cpu-baseline --no-parity --no-e2e > $O/bench_1p25e7.json 2> $O/bench_1p25e7.err
EFGP_MAX_P=128 timeout 300 python bench.py --steps 20 --warmup 5 --npts 1.25e7 --no-cpu-baseline --no-parity --no-e2e > $O/bench_1p25e7_p128.json 2> $O/bench_1p25e7_p128.err
echo "== TMA correctness"; date
EFGP_INTERP_TMA=1 timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_large_gpu.py -q -m gpu --timeout=600 > $O/pytest_tma.log 2>&1
tail -3 $O/pytest_tma.log
date
