#!/bin/bash
# round 2, GPU pass 36: ncu launch list of the default bench command's headline part on the final tree
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --suite none --no-cpu > gpurun_out/r2_36_plain.json 2> gpurun_out/r2_36_plain.err
echo "plain rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_36_launches_cfg5.csv python bench.py --steps 2 --warmup 3 --suite none --no-cpu --no-e2e > gpurun_out/r2_36_ncu.log 2>&1
echo "ncu rc=$?"
python scripts/summarize_launches.py gpurun_out/r2_36_launches_cfg5.csv | head -30
