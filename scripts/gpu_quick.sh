#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/final
mkdir -p $O
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:"efgp::|regular_fft|vector_fft|cub::" -c 900 --csv --log-file $O/ncu_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-parity --no-e2e --no-f32 > $O/ncu_launches.out 2>&1; echo "rc=$?"
tail -c 300 $O/ncu_launches.out
