# round-1 pass 29 (1 GPU): last check of the committed tree: suite, smoke, default bench
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/p29_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/p29_pytest.log
timeout 300 python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/p29_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/p29_smoke.log
timeout 600 python bench.py > gpurun_out/p29_cfg5.json 2> gpurun_out/p29_cfg5.err; echo "bench rc=$?"; tail -c 700 gpurun_out/p29_cfg5.json
