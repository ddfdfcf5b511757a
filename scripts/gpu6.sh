# round-1 pass 4: TMA window kernel parity and A/B on cfg 2
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -q --maxfail=25 --timeout 300 ) > gpurun_out/pytest_gpu.log 2>&1
tail -25 gpurun_out/pytest_gpu.log
S="python bench.py --no-cpu --no-e2e --steps 30"
run() { name=$1; shift; ( "$@" ) > gpurun_out/s_$name.json 2> gpurun_out/s_$name.err; echo "$name $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/s_$name.json | head -1)"; tail -2 gpurun_out/s_$name.err; }
run cfg2_window timeout 300 $S --workload cfg2
SLC_SPMM_WINDOW=0 run cfg2_gather timeout 300 $S --workload cfg2
timeout 300 python bench.py --no-cpu --steps 30 --workload cfg2 > gpurun_out/b_cfg2.json 2> gpurun_out/b_cfg2.err; tail -c 600 gpurun_out/b_cfg2.json
NCU="ncu --set full --clock-control none --import-source on -f"
P="python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e"
timeout 600 $NCU -k regex:k_spmm_window -c 1 -o gpurun_out/full_spmm_window_cfg2 $P --workload cfg2 > gpurun_out/full2.log 2>&1
tail -3 gpurun_out/full2.log
ls -la gpurun_out | head -40
