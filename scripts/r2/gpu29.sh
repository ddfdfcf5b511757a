#!/bin/bash
# round 2, GPU pass 29: transpose after the parallel gap fill (did the last radix pass get slower?)
mkdir -p gpurun_out
for w in cfg5t cfg3t cfg4t; do
  timeout 300 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_29_$w.json 2>/dev/null; python -c "import json;d=json.load(open('gpurun_out/r2_29_$w.json'));print('$w', d['ms_per_step'], d.get('ms_per_repeat'))"
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_29_launches_cfg5t.csv python bench.py --workload cfg5t --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_29_ncu_cfg5t.log 2>&1
python scripts/summarize_launches.py gpurun_out/r2_29_launches_cfg5t.csv | head -20
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --timeout 300 -p no:cacheprovider -k "transpose or ctrans" 2>&1 | tail -2
