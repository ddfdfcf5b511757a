#!/bin/bash
# round 2, GPU pass 4: brick SpMM (detection fixed) variants + ncu; e2e SpMV pipeline variants with a timeline
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py -m gpu -q --timeout 600 -p no:cacheprovider -k "brick or block_product" > gpurun_out/r2_04_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_04_pytest.log
tail -5 gpurun_out/r2_04_pytest.log
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 600 python bench.py --workload cfg5mm --steps 10 --warmup 3 --no-e2e > gpurun_out/r2_04_mm_$name.json 2> gpurun_out/r2_04_mm_$name.err
  echo "$name rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_04_mm_$name.json'));print(d['ms_per_step'], d['roofline']['frac'], d['parity']['ok'], d['config']['plan'][-8:])" 2>&1 | tail -1)"
}
run r2j4 SLC_BRICK_R=2 SLC_BRICK_JT=4
run r2j2 SLC_BRICK_R=2 SLC_BRICK_JT=2
run r2j8 SLC_BRICK_R=2 SLC_BRICK_JT=8
run r3j2 SLC_BRICK_R=3 SLC_BRICK_JT=2
run r3j4 SLC_BRICK_R=3 SLC_BRICK_JT=4
run r1j8 SLC_BRICK_R=1
ncu --set full --clock-control none --import-source on -k regex:k_spmm_brick -c 1 -o gpurun_out/r2_04_ncu_brick python bench.py --workload cfg5mm --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_04_ncu.log 2>&1
echo "ncu rc=$?"
e2e() {
  name=$1; shift
  env "$@" timeout 600 python bench.py --suite none --steps 20 --warmup 5 --no-cpu > gpurun_out/r2_04_e2e_$name.json 2> gpurun_out/r2_04_e2e_$name.err
  echo "e2e $name rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_04_e2e_$name.json'));print(d['ms_per_step'], d['e2e']['ms_per_step'])" 2>&1 | tail -1)"
}
e2e taper8 SLC_E2E_TAPER=1
e2e flat8 SLC_E2E_TAPER=0
e2e taper16 SLC_E2E_CHUNKS=16
e2e taper6 SLC_E2E_CHUNKS=6
e2e mode1 SLC_E2E_MODE=1
e2e nokern SLC_E2E_DIAG=4
e2e trace SLC_E2E_TRACE=1
tail -40 gpurun_out/r2_04_e2e_trace.err
