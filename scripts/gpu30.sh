# round-1 pass 28 (2 GPUs): where the all-gather SpMV spends its step (R-MAT scale 22)
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
run() { name=$1; shift; env "$@" timeout 600 $T --master-port 29541 bench.py --gpus 2 --workload cfg4 --grid 22 --steps 100 --warmup 5 > gpurun_out/p28_$name.json 2> gpurun_out/p28_$name.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/p28_$name.json").read().strip().splitlines()[-1]); print("$name", "ms=%.4f"%d["ms_per_step"], d["config"]["splits"], d["config"]["plan"][:60])
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/p28_$name.err").read()[-600:])
PY
}
run full X=1
run nopull SLC_AG_DIAG=1
run noproduct SLC_AG_DIAG=2
run nobarrier SLC_AG_DIAG=3
