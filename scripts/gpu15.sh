# round-1 pass 13 (1 GPU): ncu --set full of the SpGEMM merge kernels after the register FIFO, and of the balanced SpMV on R-MAT scale 22
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on -f"
P="python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e"
timeout 600 $NCU -k regex:k_mm_merge -c 4 -o gpurun_out/full_mm_merge_cfg3_r1d $P --workload cfg3 > gpurun_out/full_mm.log 2>&1
tail -3 gpurun_out/full_mm.log
python scripts/ncu_key.py gpurun_out/full_mm_merge_cfg3_r1d.ncu-rep > gpurun_out/r1d_ncu_full_mm_cfg3.txt; head -80 gpurun_out/r1d_ncu_full_mm_cfg3.txt
ls -la gpurun_out | head
