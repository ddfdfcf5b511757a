#!/bin/bash
# round 2, GPU pass 22 (N GPUs, N from $1): bench --gpus N with the suite after the halo staging of the block
# product; SpMV with and without programmatic dependent launch
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_22_bench_n$N.json 2> gpurun_out/r2_22_bench_n$N.err
echo "bench n$N rc=$?"
python - $N <<'PY'
import json, sys
d=json.load(open("gpurun_out/r2_22_bench_n%s.json" % sys.argv[1]))
print(d["ms_per_step"], d["value"], d.get("parity", {}).get("ok"), d["e2e"]["ms_per_step"], d["roofline"].get("boundary_wait_us"))
for e in d.get("suite",[]):
    print(" ", e.get("name"), e.get("ms_per_step"), e.get("value"), (e.get("parity") or {}).get("ok"), e.get("error"))
PY
tail -c 400 gpurun_out/r2_22_bench_n$N.err
if [ "$N" = "8" ] && [ -n "$AB" ]; then
  SLC_AG_XSPLIT=rows timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus $N --workload cfg4 --steps 20 --warmup 5 > gpurun_out/r2_22_cfg4_rows_n$N.json 2> gpurun_out/r2_22_cfg4_rows_n$N.err
  echo "cfg4 x by rows rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_22_cfg4_rows_n$N.json'));print(d['ms_per_step'])" 2>&1 | tail -1)"
  SLC_MM_PARTITION=rows timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus $N --workload cfg3 --steps 20 --warmup 5 > gpurun_out/r2_22_cfg3_rows_n$N.json 2> gpurun_out/r2_22_cfg3_rows_n$N.err
  echo "cfg3 by rows rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_22_cfg3_rows_n$N.json'));print(d['ms_per_step'])" 2>&1 | tail -1)"
fi
for pdl in 0 1; do
  SLC_PDL=$pdl timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2956$pdl bench.py --gpus $N --steps 20 --warmup 5 --suite none > gpurun_out/r2_22_pdl${pdl}_n$N.json 2> gpurun_out/r2_22_pdl${pdl}_n$N.err
  echo "pdl=$pdl rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_22_pdl${pdl}_n$N.json'));print(d['ms_per_step'], d['roofline'].get('boundary_wait_us'))" 2>&1 | tail -1)"
done
