#!/bin/bash
# Multi-GPU evidence: bash scripts/gpu_session_ngpu.sh N   (gpurun --gpus N).  Output gpurun_out/gN/.
N=${1:-2}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out/g$N
mkdir -p $O
nvidia-smi -L > $O/gpus.txt
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
if [ "$N" -ge 2 ]; then
  echo "== NCCL sharding tests"; date
  timeout 900 python -m pytest tests/test_nccl_sharding_gpu.py tests/test_probe_sharding_gpu.py -q -m gpu --timeout=600 > $O/pytest.log 2>&1
  echo "pytest rc=$?"; tail -4 $O/pytest.log
fi
port=29520
for n in $(seq 1 $N); do
  case $n in 1|2|4|8) ;; *) continue ;; esac
  port=$((port+1))
  echo "== bench $n GPUs"; date
  if [ "$n" = "1" ]; then
    timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-f32 > $O/bench_${n}gpu.json 2> $O/bench_${n}gpu.err
  else
    timeout 900 $TR --nproc-per-node $n --master-port $port bench.py --gpus $n --steps 5 --warmup 3 > $O/bench_${n}gpu.json 2> $O/bench_${n}gpu.err
  fi
  echo "rc=$?"; tail -c 300 $O/bench_${n}gpu.err
done
echo "== config4 $N GPUs"; date
timeout 900 $TR --nproc-per-node $N --master-port 29540 bench.py --workload config4 --gpus $N --steps 2 --warmup 3 > $O/bench_config4_${N}gpu.json 2> $O/bench_config4_${N}gpu.err
echo "rc=$?"; tail -c 300 $O/bench_config4_${N}gpu.err
echo "== reference arm under torchrun"; date
timeout 600 $TR --nproc-per-node $N --master-port 29541 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > $O/bench_ref_${N}gpu.json 2> $O/bench_ref_${N}gpu.err
echo "rc=$?"; head -c 300 $O/bench_ref_${N}gpu.json
python - <<PY
import json
for n in (1,2,4,8):
    try:
        d=json.loads([l for l in open("$O/bench_%dgpu.json"%n).read().splitlines() if l.startswith("{")][-1])
        print(n, "ms", round(d["ms_per_step"],3), "value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1) if d.get("e2e") else None, "parity", (d.get("parity") or {}).get("tight",{}).get("ok"))
    except Exception as e: print(n, "ERR", e)
PY
date
