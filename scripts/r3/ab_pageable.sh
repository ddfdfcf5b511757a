#!/bin/bash
# A/B of the pinned-ring staging of pageable host operands against the driver's own path (round 3)
for st in 0 1; do
  for w in cfg2 cfg5mm cfg5 cfg3 cfg3t cfg3a; do
    SLC_HOST_STAGING=$st python bench.py --workload $w --steps 20 --warmup 3 --no-cpu > gpurun_out/r3_06_ab_${w}_staging${st}.json 2>/dev/null
    python - "$w" "$st" <<'PY'
import json, sys
w, st = sys.argv[1], sys.argv[2]
d = json.loads([l for l in open("gpurun_out/r3_06_ab_%s_staging%s.json" % (w, st)) if l.startswith("{")][-1])
print(w, "staging", st, "e2e ms", round(d["e2e"]["ms_per_step"], 2), "pageable ms", (d.get("e2e_pageable") or {}).get("ms_per_step"))
PY
  done
done
