# round-1 pass 22 (1 GPU): compressed SpMV variants: 32-register cap, L2 prefetch distance
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
run() { name=$1; wl=$2; shift; shift; env "$@" timeout 900 python bench.py --workload $wl --steps 100 --warmup 10 --no-cpu --no-e2e > gpurun_out/p22_$name.json 2> gpurun_out/p22_$name.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/p22_$name.json").read().strip().splitlines()[-1])
    print("$name", "ms=%.4f"%d["ms_per_step"], "value=%.4g %s"%(d["value"], d["unit"]), "frac=%.3f"%d["roofline"]["frac"], d["roofline"].get("stored_bytes_rate_gbs"))
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/p22_$name.err").read()[-800:])
PY
}
run base cfg5 X=1
run cap32 cfg5 SLC_SELL_CAP32=1
run pf1 cfg5 SLC_SELL_PREFETCH=1
run pf2 cfg5 SLC_SELL_PREFETCH=2
run pf3 cfg5 SLC_SELL_PREFETCH=3
run cap32_pf2 cfg5 SLC_SELL_CAP32=1 SLC_SELL_PREFETCH=2
run u8_pf1 cfg5 SLC_SELL_U=8 SLC_SELL_PREFETCH=1
