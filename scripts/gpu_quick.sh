#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out/q2
timeout 900 python -m pytest tests/test_nccl_sharding_gpu.py tests/test_probe_sharding_gpu.py -q -m gpu --timeout=600 > gpurun_out/q2/pytest.log 2>&1
echo "pytest rc=$?"; grep -E "^E  |passed|failed" gpurun_out/q2/pytest.log | head
