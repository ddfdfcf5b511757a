#!/bin/bash
# round 2, GPU pass 12: assembly behind the coarse column partition, work buffers in one slab
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_properties.py -m gpu -x -q --timeout 600 -p no:cacheprovider -k "compress or triples or from_triples" > gpurun_out/r2_12_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_12_pytest.log
tail -6 gpurun_out/r2_12_pytest.log
run() {  # name, workload, env...
  name=$1; wl=$2; shift; shift
  env "$@" timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_12_$name.json 2> gpurun_out/r2_12_$name.err
  echo "$name rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_12_$name.json'));print(d['ms_per_step'], d['roofline']['frac'], d.get('parity'))" 2>&1 | tail -1)"
}
run cfg3a_coarse cfg3a SLC_ASM_COARSE=1
run cfg3a_direct cfg3a SLC_ASM_COARSE=0
run cfg3a_sort cfg3a SLC_ASM_BUCKET=0
run cfg3a_trace cfg3a SLC_ASM_TRACE=1
tail -10 gpurun_out/r2_12_cfg3a_trace.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_12_launches_cfg3a.csv python bench.py --workload cfg3a --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_12_ncu_cfg3a.log 2>&1
echo "ncu rc=$?"
python scripts/summarize_launches.py gpurun_out/r2_12_launches_cfg3a.csv | head -20
