#!/bin/bash
# round 2, GPU pass 34 (4 GPUs): bench --gpus 4 with the suite on the final tree
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29581 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/r2_34_bench_n4.json 2> gpurun_out/r2_34_bench_n4.err
echo "bench n4 rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r2_34_bench_n4.json"))
print(d["ms_per_step"], d["value"], d.get("parity", {}).get("ok"), d["e2e"]["ms_per_step"])
for e in d.get("suite",[]):
    print(" ", e.get("name"), e.get("ms_per_step"), e.get("value"), (e.get("parity") or {}).get("ok"), e.get("error"))
PY
