#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/s6
mkdir -p $O
B="--steps 5 --warmup 3 --no-cpu-baseline --no-parity --no-e2e"
for p in 512 1024 2048; do
  EFGP_MAX_P=$p timeout 300 python bench.py $B > $O/bench_p$p.json 2> $O/bench_p$p.err
done
EFGP_SIGMA=5 timeout 300 python bench.py $B > $O/bench_sigma5.json 2> $O/bench_sigma5.err
EFGP_SIGMA=5 EFGP_MAX_P=1024 timeout 300 python bench.py $B > $O/bench_sigma5_p1024.json 2> $O/bench_sigma5_p1024.err
timeout 300 python bench.py --steps 20 --warmup 5 --npts 1.25e7 --no-cpu-baseline --no-parity --no-e2e > $O/bench_1p25e7.json 2> $O/bench_1p25e7.err
date
