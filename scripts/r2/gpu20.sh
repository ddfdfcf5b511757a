#!/bin/bash
# round 2, GPU pass 20 (N GPUs, N from $1): bench --gpus N with the suite: SpGEMM by column blocks of B, power-law
# SpMV with x owned in equal slices and the pull order rotated by rank, boundary SpMM rows over column tiles
N=${1:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_multiproc.py -m gpu -q --timeout 500 -p no:cacheprovider > gpurun_out/r2_20_pytest_n$N.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_20_pytest_n$N.log
tail -4 gpurun_out/r2_20_pytest_n$N.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_20_bench_n$N.json 2> gpurun_out/r2_20_bench_n$N.err
echo "bench n$N rc=$?"
python - $N <<'PY'
import json, sys
d=json.load(open("gpurun_out/r2_20_bench_n%s.json" % sys.argv[1]))
print(d["ms_per_step"], d["value"], d.get("parity", {}).get("ok"), d["e2e"]["ms_per_step"], d["roofline"].get("boundary_wait_us"))
for e in d.get("suite",[]):
    print(" ", e.get("name"), e.get("ms_per_step"), e.get("value"), (e.get("parity") or {}).get("ok"), e.get("error"))
PY
tail -c 400 gpurun_out/r2_20_bench_n$N.err
if [ "$N" = "8" ] && [ -n "$AB" ]; then
  SLC_AG_XSPLIT=rows timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus $N --workload cfg4 --steps 20 --warmup 5 > gpurun_out/r2_20_cfg4_rows_n$N.json 2> gpurun_out/r2_20_cfg4_rows_n$N.err
  echo "cfg4 x by rows rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_20_cfg4_rows_n$N.json'));print(d['ms_per_step'])" 2>&1 | tail -1)"
  SLC_MM_PARTITION=rows timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus $N --workload cfg3 --steps 20 --warmup 5 > gpurun_out/r2_20_cfg3_rows_n$N.json 2> gpurun_out/r2_20_cfg3_rows_n$N.err
  echo "cfg3 by rows rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_20_cfg3_rows_n$N.json'));print(d['ms_per_step'])" 2>&1 | tail -1)"
fi
