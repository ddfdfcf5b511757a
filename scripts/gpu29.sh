# round-1 pass 27 (1 GPU): SpGEMM numeric merge at 5 blocks per SM (96 registers, 832-entry windows)
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
run() { name=$1; shift; env "$@" timeout 600 python bench.py --workload cfg3 --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/p27_$name.json 2> gpurun_out/p27_$name.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/p27_$name.json").read().strip().splitlines()[-1]); print("$name", "ms=%.4f"%d["ms_per_step"], "nnz/s=%.3g"%d["value"])
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/p27_$name.err").read()[-800:])
PY
}
run base X=1
run cap832 SLC_MM_STAGE_CAP=832
run minb5_cap832 SLC_MM_MINB=5 SLC_MM_STAGE_CAP=832
run minb5 SLC_MM_MINB=5
run minb6_cap832 SLC_MM_MINB=6 SLC_MM_STAGE_CAP=832
SLC_MM_MINB=5 SLC_MM_STAGE_CAP=832 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mm" > gpurun_out/p27_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/p27_pytest.log
