# round-1 pass 23 (1 GPU): compressed SpMV at 32 registers + L2 prefetch as the default: suite, default bench, distance sweep
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/p23_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/p23_pytest.log
run() { name=$1; wl=$2; shift; shift; env "$@" timeout 900 python bench.py --workload $wl --steps 100 --warmup 10 --no-cpu --no-e2e > gpurun_out/p23_$name.json 2> gpurun_out/p23_$name.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/p23_$name.json").read().strip().splitlines()[-1])
    print("$name", "ms=%.4f"%d["ms_per_step"], "value=%.4g %s"%(d["value"], d["unit"]), "frac=%.3f"%d["roofline"]["frac"], d["roofline"].get("stored_bytes_rate_gbs"))
except Exception as e:
    print("$name FAILED", e); print(open("gpurun_out/p23_$name.err").read()[-800:])
PY
}
run pf0 cfg5 SLC_SELL_PREFETCH=0
run pf1 cfg5 SLC_SELL_PREFETCH=1
run pf2 cfg5 X=1
run pf3 cfg5 SLC_SELL_PREFETCH=3
run pf5 cfg5 SLC_SELL_PREFETCH=5
timeout 600 python bench.py > gpurun_out/p23_default.json 2> gpurun_out/p23_default.err; echo "bench rc=$?"; tail -c 2700 gpurun_out/p23_default.json
NCU="ncu --set full --clock-control none --import-source on -f"
timeout 900 $NCU -k regex:k_spmv_sell -c 1 -o gpurun_out/full_spmv_sell_cfg5_r1f python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/p23_ncu.log 2>&1
python scripts/ncu_key.py gpurun_out/full_spmv_sell_cfg5_r1f.ncu-rep > gpurun_out/r1f_ncu_full_k_spmv_sell_cfg5.txt; head -34 gpurun_out/r1f_ncu_full_k_spmv_sell_cfg5.txt
