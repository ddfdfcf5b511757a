# round-1 pass 14 (1 GPU): hot-column SpMV on R-MAT (parity + A/B), SpGEMM stage window / carve-out sweep
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py -m gpu -x -q > gpurun_out/p14_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/p14_pytest.log
run() { name=$1; wl=$2; grid=$3; shift; shift; shift; env "$@" timeout 900 python bench.py --workload $wl --grid $grid --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/p14_$name.json 2> gpurun_out/p14_$name.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/p14_$name.json").read().strip().splitlines()[-1])
    print("$name", "ms=%.4f"%d["ms_per_step"], "value=%.4g %s"%(d["value"], d["unit"]), "frac=%.3f"%d["roofline"]["frac"])
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/p14_$name.err").read()[-800:])
PY
}
run mm_base cfg3 0 X=1
run mm_cap832 cfg3 0 SLC_MM_STAGE_CAP=832
run mm_cap832_c75 cfg3 0 SLC_MM_STAGE_CAP=832 SLC_MM_CARVEOUT=75
run mm_c50 cfg3 0 SLC_MM_CARVEOUT=50
run mm_c50_r2 cfg3 0 SLC_MM_CARVEOUT=50 SLC_MM_RESIDENT=2
run mm_c75_r3 cfg3 0 SLC_MM_CARVEOUT=75 SLC_MM_RESIDENT=3
run rmat22_hot cfg4 22 X=1
run rmat22_nohot cfg4 22 SLC_BAL_HOT=0
run rmat22_hot8k cfg4 22 SLC_BAL_HOT_K=8192
run rmat22_hot4k cfg4 22 SLC_BAL_HOT_K=4096
run rmat22_hot_batch cfg4 22 SLC_BAL_BATCH=1
