# round-1 pass 26 (1 GPU): validation of the final build: suite, smoke, default bench + reference arm, cfg1 / cfg2 / cfg3 / R-MAT 22 sanity lines
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/p26_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/p26_pytest.log
timeout 300 python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/p26_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/p26_smoke.log
timeout 600 python bench.py > gpurun_out/p26_cfg5.json 2> gpurun_out/p26_cfg5.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/p26_ref.json 2>&1
for wl in cfg1 cfg2 cfg3; do timeout 600 python bench.py --workload $wl --steps 50 --warmup 5 --no-cpu > gpurun_out/p26_$wl.json 2> gpurun_out/p26_$wl.err; done
timeout 600 python bench.py --workload cfg4 --grid 22 --steps 30 --warmup 3 --no-cpu --no-e2e > gpurun_out/p26_rmat22.json 2> gpurun_out/p26_rmat22.err
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/p26_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        e = d.get("e2e") or {}
        print(f.split("/")[-1], d["metric"], "ms=%.4f" % d["ms_per_step"], "value=%.5g %s" % (d["value"], d["unit"]), "frac=", (d.get("roofline") or {}).get("frac"), "e2e=", e.get("value"), e.get("ms_per_step"))
    except Exception as ex:
        print(f, "FAILED", ex)
PY
