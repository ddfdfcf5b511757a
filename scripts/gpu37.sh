# round-1 pass 35 (1 GPU): launch list of the default bench command with the final kernel
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 52 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1h_launches_cfg5.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/p35_ncu.log 2>&1
python scripts/summarize_launches.py gpurun_out/r1h_launches_cfg5.csv > gpurun_out/r1h_launches_cfg5_summary.txt; tail -6 gpurun_out/r1h_launches_cfg5_summary.txt
