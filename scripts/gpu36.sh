# round-1 pass 34 (1 GPU): step breakdown of the end-to-end SpMV
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 400 python scripts/e2e_probe.py 2>&1 | tee gpurun_out/p34_e2e_probe.txt
