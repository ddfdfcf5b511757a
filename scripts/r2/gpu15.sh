#!/bin/bash
# round 2, GPU pass 15: bucket reduce with the register sorting network; ncu --set full of the coarse scatter
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_properties.py tests/test_gpu_configs.py -m gpu -x -q --timeout 600 -p no:cacheprovider -k "compress or triples or from_triples or cfg3" > gpurun_out/r2_15_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_15_pytest.log
tail -6 gpurun_out/r2_15_pytest.log
run() {  # name, workload, env...
  name=$1; wl=$2; shift; shift
  env "$@" timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_15_$name.json 2> gpurun_out/r2_15_$name.err
  echo "$name rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_15_$name.json'));print(d['ms_per_step'], d['roofline']['frac'], d.get('parity'))" 2>&1 | tail -1)"
}
run cfg3a_coarse cfg3a SLC_ASM_COARSE=1
run cfg3a_direct cfg3a SLC_ASM_COARSE=0
run cfg3a_trace cfg3a SLC_ASM_TRACE=1
tail -10 gpurun_out/r2_15_cfg3a_trace.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_coarse_scatter -s 3 -c 1 -o gpurun_out/r2_15_ncu_coarse python bench.py --workload cfg3a --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_15_ncu_coarse.log 2>&1
echo "ncu coarse rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_bucket_reduce -s 3 -c 1 -o gpurun_out/r2_15_ncu_reduce python bench.py --workload cfg3a --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_15_ncu_reduce.log 2>&1
echo "ncu reduce rc=$?"
