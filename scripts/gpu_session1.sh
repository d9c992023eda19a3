#!/bin/bash
# Round-2 GPU session 1: parity suite, headline bench, upsampling A/B, ncu launch list + full captures.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/s1
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > $O/gpu.txt 2>&1
nproc > $O/nproc.txt
echo "== pytest" ; date
timeout 1500 python -m pytest tests -q -m gpu --timeout=900 -x --deselect tests/test_nccl_sharding_gpu.py > $O/pytest.log 2>&1
echo "pytest rc=$?" | tee -a $O/pytest.log
tail -5 $O/pytest.log
echo "== bench" ; date
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err
echo "bench rc=$?"; tail -c 1500 $O/bench.err
echo "== A/B sigma" ; date
for sg in 2 3 4 6; do
  EFGP_SIGMA=$sg timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-parity --no-e2e > $O/bench_sigma$sg.json 2> $O/bench_sigma$sg.err
done
EFGP_WIDTH_SLACK=2.5 timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-parity --no-e2e > $O/bench_slack2p5.json 2> $O/bench_slack2p5.err
echo "== reference arm" ; date
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err
echo "== ncu launch list" ; date
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/ncu_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-parity --no-e2e > $O/ncu_launches.out 2>&1
echo "== ncu full (top kernels)" ; date
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"spread2d_kernel|interp2d_kernel|bucket_scatter_kernel|bin_scatter_kernel|cg_persist_kernel|from_bucket_order|slot_to_bucket" --launch-skip 40 -c 12 -o $O/ncu_full python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-parity --no-e2e > $O/ncu_full.out 2>&1
ls -la $O
date
