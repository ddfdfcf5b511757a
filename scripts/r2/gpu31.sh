#!/bin/bash
# round 2, GPU pass 31: SpGEMM numeric merge, output staging knobs re-measured on the FIFO kernel
mkdir -p gpurun_out
run() { name=$1; shift; env "$@" timeout 300 python bench.py --workload cfg3 --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_31_$name.json 2>/dev/null; python -c "import json;d=json.load(open('gpurun_out/r2_31_$name.json'));print('$name', d['ms_per_step'], (d.get('parity') or {}).get('ok'))"; }
run default A=1
run stage0 SLC_MM_STAGE=0
run cap512 SLC_MM_STAGE_CAP=512
run cap832 SLC_MM_STAGE_CAP=832
run cap640 SLC_MM_STAGE_CAP=640
run carve SLC_MM_CARVEOUT=75
