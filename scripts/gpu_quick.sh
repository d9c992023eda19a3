#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out/q
timeout 600 python scripts/exp_sync.py > gpurun_out/q/exp_sync2.log 2>&1
grep variant gpurun_out/q/exp_sync2.log | cut -c1-200
