# round-1 pass 11 (1 GPU): e2e chunk sweep on the headline config, structural workloads for the record
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
run() { name=$1; wl=$2; shift; shift; env "$@" timeout 900 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu > gpurun_out/p11_$name.json 2> gpurun_out/p11_$name.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/p11_$name.json").read().strip().splitlines()[-1]); e=d.get("e2e") or {}
    print("$name", "ms=%.4f"%d["ms_per_step"], "value=%.4g %s"%(d["value"], d["unit"]), "frac=%.3f"%d["roofline"]["frac"], "e2e_ms=", e.get("ms_per_step"))
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/p11_$name.err").read()[-800:])
PY
}
for c in 2 4 8 16 32; do run cfg5_chunks$c cfg5 SLC_E2E_CHUNKS=$c; done
run cfg1 cfg1 X=1
run cfg3t cfg3t X=1
run cfg3a cfg3a X=1
run cfg5t cfg5t X=1
run cfg4 cfg4 X=1
run cfg4t cfg4t X=1
