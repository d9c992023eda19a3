#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out/q
timeout 900 python -m pytest tests/test_moment_spread_gpu.py -q -m gpu --timeout=600 > gpurun_out/q/pytest_mom.log 2>&1
echo "rc=$?"; grep -E "^E  |^FAILED|passed|failed" gpurun_out/q/pytest_mom.log | head -30
