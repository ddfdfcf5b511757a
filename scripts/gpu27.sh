# round-1 pass 25 (1 GPU): balanced SpMV held to 32 registers (A/B on R-MAT scale 22), suite
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/p25_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/p25_pytest.log
run() { name=$1; wl=$2; grid=$3; shift; shift; shift; env "$@" timeout 900 python bench.py --workload $wl --grid $grid --steps 30 --warmup 3 --no-cpu --no-e2e > gpurun_out/p25_$name.json 2> gpurun_out/p25_$name.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/p25_$name.json").read().strip().splitlines()[-1])
    print("$name", "ms=%.4f"%d["ms_per_step"], "value=%.4g %s"%(d["value"], d["unit"]), "frac=%.3f"%d["roofline"]["frac"])
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/p25_$name.err").read()[-800:])
PY
}
run rmat22 cfg4 22 X=1
run rmat22_minb8 cfg4 22 SLC_BAL_MINB8=1
