# round-1 pass 30 (1 GPU): smoke with the compressed kernel; ncu --set full of k_spmm_panel at its 256-thread default (cfg5mm)
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/p30_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/p30_smoke.log
NCU="ncu --set full --clock-control none --import-source on -f"
timeout 900 $NCU -k regex:k_spmm_panel -c 1 -o gpurun_out/full_spmm_panel_cfg5mm_r1g python bench.py --workload cfg5mm --steps 1 --warmup 3 --no-cpu --no-e2e > gpurun_out/p30_ncu.log 2>&1
python scripts/ncu_key.py gpurun_out/full_spmm_panel_cfg5mm_r1g.ncu-rep > gpurun_out/r1g_ncu_full_spmm_panel_cfg5mm.txt; head -36 gpurun_out/r1g_ncu_full_spmm_panel_cfg5mm.txt
