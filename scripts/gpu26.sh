# round-1 pass 24 (8 GPUs): strong scaling of the headline SpMV at N = 1, 2, 4, 8 and the other row-partitioned workloads at N = 8
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
T() { n=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) bench.py --gpus $n "$@"; }
timeout 600 python bench.py --steps 200 --warmup 10 --no-cpu > gpurun_out/p24_n1_spmv.json 2> gpurun_out/p24_n1_spmv.err
for n in 2 4 8; do
  timeout 600 bash -c "$(declare -f T); T $n --steps 200 --warmup 10" > gpurun_out/p24_n${n}_spmv.json 2> gpurun_out/p24_n${n}_spmv.err
done
timeout 900 bash -c "$(declare -f T); T 8 --workload cfg5mm --steps 20 --warmup 3" > gpurun_out/p24_n8_spmm.json 2> gpurun_out/p24_n8_spmm.err
timeout 600 bash -c "$(declare -f T); T 8 --workload cfg3 --steps 10 --warmup 3" > gpurun_out/p24_n8_cfg3.json 2> gpurun_out/p24_n8_cfg3.err
timeout 900 bash -c "$(declare -f T); T 8 --workload cfg4 --grid 22 --steps 50 --warmup 5" > gpurun_out/p24_n8_cfg4.json 2> gpurun_out/p24_n8_cfg4.err
timeout 600 bash -c "$(declare -f T); T 8 --scaling weak --steps 100 --warmup 10 --no-e2e" > gpurun_out/p24_n8_spmv_weak.json 2> gpurun_out/p24_n8_spmv_weak.err
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/p24_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        e = d.get("e2e") or {}
        print(f, d["metric"], "n=%d" % d["n_gpus"], "ms=%.4f" % d["ms_per_step"], "value=%.5g %s" % (d["value"], d["unit"]), "e2e=", e.get("value"))
    except Exception as ex:
        print(f, "FAILED", ex); print(open(f.replace(".json", ".err")).read()[-600:])
PY
