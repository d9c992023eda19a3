"""Print the per-stage table of bench.py JSON lines (one file per argument)."""
import json, sys
for f in sys.argv[1:]:
    for line in open(f):
        line = line.strip()
        if not line.startswith("{"):
            continue
        d = json.loads(line)
        e = d.get("e2e") or {}
        print(f"== {f}: value {d['value']:.0f} {d['unit']}  {d['ms_per_step']:.3f} ms/step  e2e {e.get('value')}  launches {d.get('gpu_launches')}")
        tot = 0
        for k, v in (d.get("stages") or {}).items():
            tot += v["ms_per_step"]
            print(f"   {k:28s} {v['ms_per_step']:8.3f} ms  x{v['launches_per_step']:.0f}  hbm {v.get('hbm_frac', 0):.3f}")
        print(f"   sum of stages {tot:.3f} ms")
