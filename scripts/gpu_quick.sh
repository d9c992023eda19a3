#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/q
mkdir -p $O
for i in 1 2 3 4; do
  timeout 900 python -m pytest tests -x -q -m gpu --timeout=600 > $O/pytest_rep$i.log 2>&1
  echo "run $i rc=$? $(tail -1 $O/pytest_rep$i.log)"
  grep -E "^FAILED|^E  " $O/pytest_rep$i.log | head -5
done
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
