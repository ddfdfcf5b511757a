#!/bin/bash
# round 2, GPU pass 32: whole GPU suite, smoke(), default bench line and the reference arm on the final tree
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 -p no:cacheprovider > gpurun_out/r2_32_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_32_pytest.log
tail -5 gpurun_out/r2_32_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_32_bench.json 2> gpurun_out/r2_32_bench.err
echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r2_32_bench.json"))
print(d["ms_per_step"], d["value"], d["roofline"]["frac"], d["e2e"]["ms_per_step"], d.get("parity",{}).get("ok"), d.get("gpu_launches"))
for e in d.get("suite",[]):
    print(" ", e.get("name"), e.get("ms_per_step"), (e.get("roofline") or {}).get("frac"), (e.get("parity") or {}).get("ok"), e.get("error"))
PY
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_32_ref.json 2> gpurun_out/r2_32_ref.err
echo "ref rc=$? $(cut -c1-300 gpurun_out/r2_32_ref.json)"
