# round-1 pass 19 (1 GPU): ncu of the headline kernel with compressed indices (launch list + full set), and of the balanced SpMV with the packed hot table
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/p19_plain.json 2> gpurun_out/p19_plain.err; tail -c 400 gpurun_out/p19_plain.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1e_launches_cfg5.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/p19_ncu1.log 2>&1
python scripts/summarize_launches.py gpurun_out/r1e_launches_cfg5.csv > gpurun_out/r1e_launches_cfg5_summary.txt; cat gpurun_out/r1e_launches_cfg5_summary.txt
NCU="ncu --set full --clock-control none --import-source on -f"
timeout 900 $NCU -k regex:k_spmv_sell -c 1 -o gpurun_out/full_spmv_sell_cfg5_r1e python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/p19_ncu2.log 2>&1
python scripts/ncu_key.py gpurun_out/full_spmv_sell_cfg5_r1e.ncu-rep > gpurun_out/r1e_ncu_full_k_spmv_sell_cfg5.txt; cat gpurun_out/r1e_ncu_full_k_spmv_sell_cfg5.txt
timeout 900 $NCU -k regex:k_spmv_bal -c 2 -o gpurun_out/full_spmv_bal_rmat22_r1e python bench.py --workload cfg4 --grid 22 --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/p19_ncu3.log 2>&1
python scripts/ncu_key.py gpurun_out/full_spmv_bal_rmat22_r1e.ncu-rep > gpurun_out/r1e_ncu_full_k_spmv_bal_rmat22.txt; head -40 gpurun_out/r1e_ncu_full_k_spmv_bal_rmat22.txt
