# round-1 pass 21 (1 GPU): final form of the compressed-index SpMV: full suite, smoke, default bench, cfg1, ncu captures
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/p21_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/p21_pytest.log
timeout 300 python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/p21_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/p21_smoke.log
timeout 600 python bench.py > gpurun_out/p21_cfg5.json 2> gpurun_out/p21_cfg5.err; echo "bench rc=$?"; tail -c 2700 gpurun_out/p21_cfg5.json
timeout 600 python bench.py --workload cfg1 --no-cpu > gpurun_out/p21_cfg1.json 2> gpurun_out/p21_cfg1.err; tail -c 900 gpurun_out/p21_cfg1.json
bash scripts/gpu21.sh
