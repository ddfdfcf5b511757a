# round-1 pass 17 (1 GPU): SELL index compression (one base per (slice, k)) parity and A/B on cfg5 / cfg1
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/p17_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/p17_pytest.log
run() { name=$1; wl=$2; shift; shift; env "$@" timeout 900 python bench.py --workload $wl --steps 100 --warmup 10 --no-cpu --no-e2e > gpurun_out/p17_$name.json 2> gpurun_out/p17_$name.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/p17_$name.json").read().strip().splitlines()[-1])
    print("$name", "ms=%.4f"%d["ms_per_step"], "value=%.4g %s"%(d["value"], d["unit"]), "frac=%.3f"%d["roofline"]["frac"], d["roofline"].get("stored_bytes_rate_gbs"), d["config"]["plan"])
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/p17_$name.err").read()[-800:])
PY
}
run cfg5_off cfg5 SLC_SELL_KBASE=0
run cfg5_kb_u8 cfg5 X=1
run cfg5_kb_u4 cfg5 SLC_SELL_U=4
run cfg5_kb_u16 cfg5 SLC_SELL_U=16
run cfg1_off cfg1 SLC_SELL_KBASE=0
run cfg1_kb cfg1 X=1
