#!/bin/bash
# Round-2 evidence session (1 GPU): full GPU suite, headline bench with parity / CPU arm / e2e / f32, shard-size and
# config-4 figures, the reference arm, then the ncu launch list and one --set full capture of the hot kernels.
# Output: gpurun_out/final/ -> python scripts/make_profiles_r02.py gpurun_out/final
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/final
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/gpu.txt 2>&1
echo "== pytest"; date
timeout 2400 python -m pytest tests -q -m gpu --timeout=900 > $O/pytest.log 2>&1
echo "pytest rc=$?"; tail -2 $O/pytest.log
timeout 300 python __graft_entry__.py smoke > $O/smoke.log 2>&1; tail -1 $O/smoke.log
echo "== bench (default)"; date
timeout 1500 python bench.py > $O/bench.json 2> $O/bench.err; echo "rc=$?"
echo "== bench shard size"; date
timeout 600 python bench.py --steps 20 --warmup 5 --npts 1.25e7 --no-cpu-baseline --no-parity > $O/bench_shard.json 2> $O/bench_shard.err; echo "rc=$?"
echo "== bench presort"; date
timeout 600 python bench.py --presort --no-cpu-baseline --no-parity --no-e2e --no-f32 > $O/bench_presort.json 2> $O/bench_presort.err; echo "rc=$?"
echo "== bench config4"; date
timeout 900 python bench.py --workload config4 --steps 3 --warmup 3 > $O/bench_config4.json 2> $O/bench_config4.err; echo "rc=$?"
echo "== reference arm"; date
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "rc=$?"
echo "== ncu launch list"; date
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:"efgp::|regular_fft|vector_fft|cub::" -c 900 --csv --log-file $O/ncu_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-parity --no-e2e --no-f32 > $O/ncu_launches.out 2>&1; echo "rc=$?"
echo "== ncu full"; date
# (the .ncu-rep with sources is > 64 MiB, more than gpurun copies back: keep it on the box, bring the CSV pages)
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"spread2d_|moment_to_grid|interp2d_kernel|bucket_|bin_|cg_conv|cg_persist|from_bucket|slot_to_bucket|bbox_partial" --launch-skip 36 -c 14 -o /tmp/ncu_full python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-parity --no-e2e --no-f32 > $O/ncu_full.out 2>&1; echo "rc=$?"
ncu -i /tmp/ncu_full.ncu-rep --page raw --csv > $O/ncu_full_raw.csv 2>/dev/null
ncu -i /tmp/ncu_full.ncu-rep --page details --csv > $O/ncu_full_details.csv 2>/dev/null
ls -la $O; date
