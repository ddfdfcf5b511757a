#!/bin/bash
# round 2, GPU pass 35: assembly: coarse scatter with packed registers (64, 4 blocks per SM), 63-exchange sorting network
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_properties.py tests/test_gpu_configs.py -m gpu -x -q --timeout 300 -p no:cacheprovider -k "compress or triples or cfg3" > gpurun_out/r2_35_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_35_pytest.log
tail -3 gpurun_out/r2_35_pytest.log
for i in 1 2; do
timeout 300 python bench.py --workload cfg3a --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_35_cfg3a_$i.json 2>/dev/null; python -c "import json;d=json.load(open('gpurun_out/r2_35_cfg3a_$i.json'));print('cfg3a', d['ms_per_step'])"
done
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_35_launches_cfg3a.csv python bench.py --workload cfg3a --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_35_ncu.log 2>&1
python scripts/summarize_launches.py gpurun_out/r2_35_launches_cfg3a.csv | head -14
