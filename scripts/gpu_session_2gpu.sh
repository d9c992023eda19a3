#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/g2
mkdir -p $O
nvidia-smi -L > $O/gpus.txt
echo "== NCCL sharding tests"; date
timeout 900 python -m pytest tests/test_nccl_sharding_gpu.py tests/test_probe_sharding_gpu.py -q -m gpu --timeout=600 > $O/pytest.log 2>&1
echo "pytest rc=$?"; tail -8 $O/pytest.log
echo "== bench 2 GPUs"; date
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > $O/bench_2gpu.json 2> $O/bench_2gpu.err
echo "rc=$?"; tail -c 600 $O/bench_2gpu.err
echo "== config4 2 GPUs"; date
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --workload config4 --gpus 2 --steps 2 --warmup 3 > $O/bench_config4_2gpu.json 2> $O/bench_config4_2gpu.err
echo "rc=$?"; tail -c 600 $O/bench_config4_2gpu.err
echo "== reference arm under torchrun"; date
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > $O/bench_ref_2gpu.json 2> $O/bench_ref_2gpu.err
echo "rc=$?"; head -c 300 $O/bench_ref_2gpu.json
date
