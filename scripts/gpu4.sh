# round-1 pass 2: parity of the new kernels, bench lines for every workload, ncu of each new kernel
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -q --maxfail=25 --timeout 300 ) > gpurun_out/pytest_gpu.log 2>&1
tail -15 gpurun_out/pytest_gpu.log
B="python bench.py --no-cpu"
for w in cfg2 cfg5mm cfg3 cfg3t cfg3a cfg5t cfg1; do
  timeout 600 $B --workload $w --steps 30 > gpurun_out/b_$w.json 2> gpurun_out/b_$w.err; tail -c 1200 gpurun_out/b_$w.json; tail -2 gpurun_out/b_$w.err
done
timeout 600 $B --workload cfg5 --steps 100 > gpurun_out/b_cfg5.json 2> gpurun_out/b_cfg5.err; tail -c 1500 gpurun_out/b_cfg5.json
# sweeps (device-only numbers)
S="$B --no-e2e --steps 20"
SLC_PANEL_THREADS=256 timeout 300 $S --workload cfg5mm > gpurun_out/s_mm_t256.json 2>&1
SLC_PANEL_THREADS=256 SLC_PANEL_SMEM_KB=48 timeout 300 $S --workload cfg5mm > gpurun_out/s_mm_t256_s48.json 2>&1
SLC_PANEL_THREADS=64 timeout 300 $S --workload cfg5mm > gpurun_out/s_mm_t64.json 2>&1
SLC_PANEL_JT=4 timeout 300 $S --workload cfg5mm > gpurun_out/s_mm_jt4.json 2>&1
SLC_PANEL_KU=4 timeout 300 $S --workload cfg5mm > gpurun_out/s_mm_ku4.json 2>&1
SLC_SPMM_PANEL=0 timeout 300 $S --workload cfg5mm > gpurun_out/s_mm_old.json 2>&1
SLC_SPMM_STRIDE=1024 timeout 300 $S --workload cfg2 > gpurun_out/s_cfg2_stride.json 2>&1
SLC_SPMM_THREADS=256 timeout 300 $S --workload cfg2 > gpurun_out/s_cfg2_t256.json 2>&1
for f in gpurun_out/s_*.json; do echo $f; grep -o '"ms_per_step": [0-9.]*' $f | head -1; done
NCU="ncu --set full --clock-control none --import-source on -f"
P="python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e"
timeout 600 $NCU -k regex:k_spmm_panel -c 1 -o gpurun_out/full_spmm_panel_cfg5mm $P --workload cfg5mm > gpurun_out/full1.log 2>&1
timeout 600 $NCU -k regex:k_spmm_short -c 1 -o gpurun_out/full_spmm_short_cfg2 $P --workload cfg2 > gpurun_out/full2.log 2>&1
timeout 600 $NCU -k regex:k_mm_merge -c 2 -o gpurun_out/full_mm_cfg3 $P --workload cfg3 > gpurun_out/full3.log 2>&1
timeout 600 $NCU -k regex:"k_radix_scatter|k_radix_hist|k_transpose_gather|k_expand_cols" -s 6 -c 6 -o gpurun_out/full_transpose_cfg5t $P --workload cfg5t > gpurun_out/full4.log 2>&1
for w in cfg5mm cfg2 cfg3 cfg5t cfg3a; do
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$w.csv python bench.py --workload $w --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_$w.log 2>&1
done
( time timeout 900 $B --workload cfg4 --steps 30 ) > gpurun_out/b_cfg4.json 2> gpurun_out/b_cfg4.err; tail -c 1500 gpurun_out/b_cfg4.json; tail -3 gpurun_out/b_cfg4.err
timeout 900 $NCU -k regex:k_spmv_merge -c 1 -o gpurun_out/full_merge_cfg4 $P --workload cfg4 > gpurun_out/full5.log 2>&1
ls -la gpurun_out
