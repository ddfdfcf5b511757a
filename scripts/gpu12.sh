# round-1 pass 10 (1 GPU): SpGEMM merge with a two-deep register FIFO per list
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py tests/test_gpu_properties.py -m gpu -x -q > gpurun_out/p10_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/p10_pytest.log
B="timeout 600 python bench.py --workload cfg3 --steps 10 --warmup 3 --no-cpu --no-e2e"
run() { name=$1; shift; env "$@" $B > gpurun_out/p10_$name.json 2> gpurun_out/p10_$name.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/p10_$name.json").read().strip().splitlines()[-1]); print("$name", "ms=%.4f"%d["ms_per_step"], "nnz/s=%.3g"%d["value"])
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/p10_$name.err").read()[-800:])
PY
}
run base X=1
run nostage SLC_MM_STAGE=0
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/p10_launches_cfg3.csv python bench.py --workload cfg3 --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/p10_ncu.log 2>&1
python scripts/summarize_launches.py gpurun_out/p10_launches_cfg3.csv | tail -25
