#!/bin/bash
# round 2, GPU pass 33: SpGEMM output window: fitted to the product (auto) against fixed sizes, same box
mkdir -p gpurun_out
run() { name=$1; shift; env "$@" timeout 300 python bench.py --workload cfg3 --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_33_$name.json 2>/dev/null; python -c "import json;d=json.load(open('gpurun_out/r2_33_$name.json'));print('$name', d['ms_per_step'], d.get('ms_per_repeat'))"; }
run auto A=1
run cap1024 SLC_MM_STAGE_CAP=1024
run cap832 SLC_MM_STAGE_CAP=832
run cap800 SLC_MM_STAGE_CAP=800
run auto_b A=1
run cap1024_b SLC_MM_STAGE_CAP=1024
