#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/q
mkdir -p $O
for i in 1 2 3 4 5 6 7 8; do
  timeout 900 python -m pytest tests -q -m gpu --timeout=600 > $O/pytest_rep$i.log 2>&1
  echo "run $i rc=$? $(tail -1 $O/pytest_rep$i.log)"
  grep -E "^FAILED" $O/pytest_rep$i.log | head -5
done
