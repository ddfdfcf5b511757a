# round-1 pass 8 (1 GPU): k_spmm_panel line tiling A/B on cfg5mm
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py -m gpu -x -q -k "mulm or block" > gpurun_out/p8_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/p8_pytest.log
B="timeout 600 python bench.py --workload cfg5mm --steps 10 --warmup 3 --no-cpu --no-e2e"
run() { name=$1; shift; env "$@" $B > gpurun_out/p8_$name.json 2> gpurun_out/p8_$name.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/p8_$name.json").read().strip().splitlines()[-1]); print("$name", "ms=%.3f"%d["ms_per_step"], "GB/s=%.0f"%d["value"])
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/p8_$name.err").read()[-800:])
PY
}
run consecutive SLC_PANEL_STRIDE=0
run tiled X=1
run tiled_w2 SLC_PANEL_WIDTH=2
run tiled_w4 SLC_PANEL_WIDTH=4
run tiled_512 SLC_PANEL_THREADS=512
run tiled_512_w2 SLC_PANEL_THREADS=512 SLC_PANEL_WIDTH=2
run tiled_384 SLC_PANEL_THREADS=384
run tiled_jt4 SLC_PANEL_JT=4
