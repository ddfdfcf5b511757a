# round-1 pass 15 (1 GPU): hot-column table variants on R-MAT scale 22
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
SLC_BAL_HOT_MODE=2 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "hot or mulv" > gpurun_out/p15_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/p15_pytest.log
SLC_BAL_HOT_GROUPS=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "hot" > gpurun_out/p15_pytest_g1.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/p15_pytest_g1.log
run() { name=$1; wl=$2; grid=$3; shift; shift; shift; env "$@" timeout 900 python bench.py --workload $wl --grid $grid --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/p15_$name.json 2> gpurun_out/p15_$name.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/p15_$name.json").read().strip().splitlines()[-1])
    print("$name", "ms=%.4f"%d["ms_per_step"], "value=%.4g %s"%(d["value"], d["unit"]), "frac=%.3f"%d["roofline"]["frac"])
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/p15_$name.err").read()[-800:])
PY
}
run nohot cfg4 22 SLC_BAL_HOT=0
for k in 16384 8192 4096 2048 512; do run packed_k$k cfg4 22 SLC_BAL_HOT_MODE=2 SLC_BAL_HOT_K=$k; done
run smem_g2_k4096 cfg4 22 SLC_BAL_HOT_GROUPS=2 SLC_BAL_HOT_K=4096
run smem_g1_k4096 cfg4 22 SLC_BAL_HOT_GROUPS=1 SLC_BAL_HOT_K=4096
run smem_g1_k2048 cfg4 22 SLC_BAL_HOT_GROUPS=1 SLC_BAL_HOT_K=2048
run smem_g2_k12288 cfg4 22 SLC_BAL_HOT_GROUPS=2 SLC_BAL_HOT_K=12288
