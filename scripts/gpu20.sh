# round-1 pass 18 (1 GPU): index compression with U=4, full GPU suite, cfg1 A/B, 2-D check
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/p18_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/p18_pytest.log
run() { name=$1; wl=$2; shift; shift; env "$@" timeout 900 python bench.py --workload $wl --steps 100 --warmup 10 --no-cpu --no-e2e > gpurun_out/p18_$name.json 2> gpurun_out/p18_$name.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/p18_$name.json").read().strip().splitlines()[-1])
    print("$name", "ms=%.4f"%d["ms_per_step"], "value=%.4g %s"%(d["value"], d["unit"]), "frac=%.3f"%d["roofline"]["frac"], d["roofline"].get("stored_bytes_rate_gbs"))
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/p18_$name.err").read()[-800:])
PY
}
run cfg5_kb cfg5 X=1
run cfg1_off cfg1 SLC_SELL_KBASE=0
run cfg1_kb cfg1 X=1
run cfg1_off_u4 cfg1 SLC_SELL_KBASE=0 SLC_SELL_U=4
timeout 600 python bench.py --steps 200 --warmup 10 > gpurun_out/p18_default.json 2> gpurun_out/p18_default.err; tail -c 2600 gpurun_out/p18_default.json
