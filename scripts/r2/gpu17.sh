#!/bin/bash
# round 2, GPU pass 17: whole GPU suite on the current tree, default bench line (headline + suite),
# reference arm, ncu of the coarse scatter without atomics
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 -p no:cacheprovider > gpurun_out/r2_17_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_17_pytest.log
tail -5 gpurun_out/r2_17_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_17_bench.json 2> gpurun_out/r2_17_bench.err
echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r2_17_bench.json"))
print(d["ms_per_step"], d["value"], d["roofline"]["frac"], d["e2e"]["ms_per_step"], d.get("parity"))
for e in d.get("suite",[]):
    print(" ", e.get("name"), e.get("ms_per_step"), (e.get("roofline") or {}).get("frac"), (e.get("parity") or {}).get("ok"), e.get("error"))
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_coarse_scatter -s 3 -c 1 -o gpurun_out/r2_17_ncu_coarse python bench.py --workload cfg3a --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_17_ncu_coarse.log 2>&1
echo "ncu coarse rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_coo_check_count -s 3 -c 1 -o gpurun_out/r2_17_ncu_count python bench.py --workload cfg3a --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_17_ncu_count.log 2>&1
echo "ncu count rc=$?"
