#!/bin/bash
# round 2, GPU pass 3: brick SpMM parity + variants on cfg5mm, ncu of the winner
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py -m gpu -q --timeout 600 -p no:cacheprovider -k "brick or block_product" > gpurun_out/r2_03_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_03_pytest.log
tail -15 gpurun_out/r2_03_pytest.log
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 600 python bench.py --workload cfg5mm --steps 10 --warmup 3 --no-e2e > gpurun_out/r2_03_mm_$name.json 2> gpurun_out/r2_03_mm_$name.err
  echo "$name rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_03_mm_$name.json'));print(d['ms_per_step'], d['roofline']['frac'], d['parity']['ok'], d['config']['plan'][-8:])" 2>&1 | tail -1)"
}
run r2j4 SLC_BRICK_R=2 SLC_BRICK_JT=4
run r2j2 SLC_BRICK_R=2 SLC_BRICK_JT=2
run r2j8 SLC_BRICK_R=2 SLC_BRICK_JT=8
run r3j2 SLC_BRICK_R=3 SLC_BRICK_JT=2
run r3j4 SLC_BRICK_R=3 SLC_BRICK_JT=4
run r1j8 SLC_BRICK_R=1
run panel SLC_SPMM_BRICK=0
ncu --set full --clock-control none --import-source on -k regex:k_spmm_brick -c 1 -o gpurun_out/r2_03_ncu_brick python bench.py --workload cfg5mm --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_03_ncu.log 2>&1
echo "ncu rc=$?"
