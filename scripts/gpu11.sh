# round-1 pass 9 (1 GPU): SpGEMM numeric merge with a bounded number of resident blocks (L1 working set of A)
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -x -q -k "mm or panel or spgemm" > gpurun_out/p9_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/p9_pytest.log
SLC_MM_RESIDENT=3 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mm" > gpurun_out/p9_pytest_res.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/p9_pytest_res.log
B="timeout 600 python bench.py --workload cfg3 --steps 10 --warmup 3 --no-cpu --no-e2e"
run() { name=$1; shift; env "$@" $B > gpurun_out/p9_$name.json 2> gpurun_out/p9_$name.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/p9_$name.json").read().strip().splitlines()[-1]); print("$name", "ms=%.4f"%d["ms_per_step"], "nnz/s=%.3g"%d["value"])
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/p9_$name.err").read()[-800:])
PY
}
run base X=1
for r in 1 2 3 4 6 8 12; do run res$r SLC_MM_RESIDENT=$r; done
run res2_nostage SLC_MM_RESIDENT=2 SLC_MM_STAGE=0
run res4_nostage SLC_MM_RESIDENT=4 SLC_MM_STAGE=0
run res8_nostage SLC_MM_RESIDENT=8 SLC_MM_STAGE=0
