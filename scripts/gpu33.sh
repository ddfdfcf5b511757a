# round-1 pass 31 (1 GPU): ncu --set full of the balanced SpMV with the packed hot table at full size (R-MAT scale 24)
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on -f"
timeout 1200 $NCU -k regex:"k_spmv_bal|k_hot_pack" -c 3 -o gpurun_out/full_spmv_bal_cfg4_r1g python bench.py --workload cfg4 --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/p31_ncu.log 2>&1
tail -3 gpurun_out/p31_ncu.log
python scripts/ncu_key.py gpurun_out/full_spmv_bal_cfg4_r1g.ncu-rep > gpurun_out/r1g_ncu_full_k_spmv_bal_cfg4.txt; grep -E "====|gpu__time|dram__bytes|hit_rate|registers" gpurun_out/r1g_ncu_full_k_spmv_bal_cfg4.txt
