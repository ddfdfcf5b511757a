# round-1 pass 3: parity, then A/B of the variants added after pass 2
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -q --maxfail=25 --timeout 300 ) > gpurun_out/pytest_gpu.log 2>&1
tail -15 gpurun_out/pytest_gpu.log
S="python bench.py --no-cpu --no-e2e --steps 20"
run() { name=$1; shift; ( "$@" ) > gpurun_out/s_$name.json 2> gpurun_out/s_$name.err; echo "$name $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/s_$name.json | head -1)"; }
run cfg2_pipe1 timeout 300 $S --workload cfg2
SLC_SPMM_PIPE=0 run cfg2_pipe0 timeout 300 $S --workload cfg2
run mm_t256 timeout 300 $S --workload cfg5mm
SLC_PANEL_THREADS=512 run mm_t512 timeout 300 $S --workload cfg5mm
SLC_PANEL_THREADS=512 SLC_PANEL_KU=4 run mm_t512_ku4 timeout 300 $S --workload cfg5mm
SLC_PANEL_KU=4 run mm_t256_ku4 timeout 300 $S --workload cfg5mm
SLC_PANEL_THREADS=384 run mm_t384 timeout 300 $S --workload cfg5mm
run cfg3_stage1 timeout 300 $S --workload cfg3
SLC_MM_STAGE=0 run cfg3_stage0 timeout 300 $S --workload cfg3
run cfg3t timeout 300 $S --workload cfg3t
run cfg5t timeout 300 $S --workload cfg5t
run cfg3a timeout 300 $S --workload cfg3a
run cfg4_smem timeout 900 $S --workload cfg4
SLC_BAL_BATCH=1 run cfg4_batch timeout 900 $S --workload cfg4
NCU="ncu --set full --clock-control none --import-source on -f"
P="python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e"
timeout 600 $NCU -k regex:k_mm_merge -s 2 -c 2 -o gpurun_out/full_mm_numeric_cfg3 $P --workload cfg3 > gpurun_out/full3.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_cfg3.csv python bench.py --workload cfg3 --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_cfg3.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_cfg5t.csv python bench.py --workload cfg5t --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_cfg5t.log 2>&1
ls -la gpurun_out | head -50
