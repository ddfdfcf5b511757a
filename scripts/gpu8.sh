# round-1 pass 6 (1 GPU): full GPU test suite, default bench + reference arm, SpMM end-to-end with the column-group pipeline
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/p6_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/p6_pytest.log
timeout 300 python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/p6_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/p6_smoke.log
timeout 600 python bench.py > gpurun_out/p6_cfg5.json 2> gpurun_out/p6_cfg5.err; echo "bench rc=$?"; tail -c 2500 gpurun_out/p6_cfg5.json
timeout 300 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/p6_ref.json 2>&1; tail -c 600 gpurun_out/p6_ref.json
timeout 600 python bench.py --workload cfg2 --steps 50 --warmup 5 --no-cpu > gpurun_out/p6_cfg2.json 2> gpurun_out/p6_cfg2.err; tail -c 1500 gpurun_out/p6_cfg2.json
SLC_E2E_COLGROUPS=1 timeout 600 python bench.py --workload cfg2 --steps 10 --warmup 3 --no-cpu > gpurun_out/p6_cfg2_nopipe.json 2> gpurun_out/p6_cfg2_nopipe.err; tail -c 700 gpurun_out/p6_cfg2_nopipe.json
timeout 600 python bench.py --workload cfg5mm --steps 10 --warmup 3 --no-cpu > gpurun_out/p6_cfg5mm.json 2> gpurun_out/p6_cfg5mm.err; tail -c 1500 gpurun_out/p6_cfg5mm.json
