# diagnostic pass: GPU tests, then one ncu --set full capture of each weak kernel
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1
tail -5 gpurun_out/pytest_gpu.log
NCU="ncu --set full --clock-control none --import-source on -f"
B="python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e"
timeout 600 $NCU -k regex:k_spmm_rows -c 1 -o gpurun_out/full_spmm_rows_cfg5mm $B --workload cfg5mm > gpurun_out/full1.log 2>&1
timeout 600 $NCU -k regex:k_spmm_short -c 1 -o gpurun_out/full_spmm_short_cfg2 $B --workload cfg2 > gpurun_out/full2.log 2>&1
timeout 600 $NCU -k regex:"k_mm_(numeric|symbolic)" -c 2 -o gpurun_out/full_mm_cfg3 $B --workload cfg3 > gpurun_out/full3.log 2>&1
timeout 600 $NCU -k regex:"k_radix_scatter|k_transpose_gather|k_radix_hist" -c 3 -o gpurun_out/full_radix_cfg5 $B --workload cfg5 > gpurun_out/full4.log 2>&1
timeout 900 $NCU -k regex:"k_spmv_bal" -c 1 -o gpurun_out/full_bal_cfg4 $B --workload cfg4 > gpurun_out/full5.log 2>&1
tail -3 gpurun_out/full*.log
ls -la gpurun_out
