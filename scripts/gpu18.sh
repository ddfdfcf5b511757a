# round-1 pass 16 (1 GPU): hot-column table at full size (R-MAT scale 24), packed vs off
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
run() { name=$1; wl=$2; grid=$3; shift; shift; shift; env "$@" timeout 900 python bench.py --workload $wl --grid $grid --steps 30 --warmup 3 --no-cpu --no-e2e > gpurun_out/p16_$name.json 2> gpurun_out/p16_$name.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/p16_$name.json").read().strip().splitlines()[-1])
    print("$name", "ms=%.4f"%d["ms_per_step"], "value=%.4g %s"%(d["value"], d["unit"]), "frac=%.3f"%d["roofline"]["frac"])
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/p16_$name.err").read()[-800:])
PY
}
run nohot cfg4 24 SLC_BAL_HOT=0
run packed_k8192 cfg4 24 SLC_BAL_HOT_MODE=2 SLC_BAL_HOT_K=8192
run packed_k16384 cfg4 24 SLC_BAL_HOT_MODE=2
