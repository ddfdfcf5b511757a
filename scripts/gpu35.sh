# round-1 pass 33 (1 GPU): PCIe floor of the end-to-end SpMV line
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python scripts/pcie_probe.py | tee gpurun_out/p33_pcie.txt
