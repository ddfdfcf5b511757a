#!/bin/bash
# round 2, GPU pass 19 (2 GPUs): multi-process / multi-device tests on the current tree; bench --gpus 2 with the
# suite (boundary rows of the block product now spread over column tiles)
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2_19_gpus.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_multiproc.py tests/test_gpu_boundary.py tests/test_c_abi.py tests/test_gpu_sharded.py -m gpu -q --timeout 600 -p no:cacheprovider > gpurun_out/r2_19_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_19_pytest.log
tail -5 gpurun_out/r2_19_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_19_bench_n2.json 2> gpurun_out/r2_19_bench_n2.err
echo "bench n2 rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r2_19_bench_n2.json"))
print(d["ms_per_step"], d["value"], d.get("parity"), d["e2e"]["ms_per_step"])
for e in d.get("suite",[]):
    print(" ", e.get("name"), e.get("ms_per_step"), (e.get("parity") or {}).get("ok"), e.get("error"))
PY
tail -c 600 gpurun_out/r2_19_bench_n2.err
