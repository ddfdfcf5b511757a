#!/bin/bash
# round 2, GPU pass 30: SpGEMM pointer pass (32-bit scan + 64-bit total instead of widen / 64-bit scan / narrow)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_properties.py tests/test_gpu_sharded.py -m gpu -x -q --timeout 300 -p no:cacheprovider -k "mm or mul_laws or spgemm" > gpurun_out/r2_30_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_30_pytest.log
tail -3 gpurun_out/r2_30_pytest.log
for i in 1 2; do
timeout 300 python bench.py --workload cfg3 --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_30_cfg3_$i.json 2>/dev/null; python -c "import json;d=json.load(open('gpurun_out/r2_30_cfg3_$i.json'));print('cfg3', d['ms_per_step'], d.get('parity'))"
done
