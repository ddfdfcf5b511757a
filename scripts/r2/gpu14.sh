#!/bin/bash
# round 2, GPU pass 14: staged coarse partition for the assembly; clock sampler at 25 ms (was 2 ms:
# its NVML queries stalled the host synchronisations of the timed step)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_properties.py -m gpu -x -q --timeout 600 -p no:cacheprovider -k "compress or triples or from_triples" > gpurun_out/r2_14_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_14_pytest.log
tail -6 gpurun_out/r2_14_pytest.log
run() {  # name, workload, env...
  name=$1; wl=$2; shift; shift
  env "$@" timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_14_$name.json 2> gpurun_out/r2_14_$name.err
  echo "$name rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_14_$name.json'));print(d['ms_per_step'], d['roofline']['frac'], d.get('parity'))" 2>&1 | tail -1)"
}
run cfg3a_coarse cfg3a SLC_ASM_COARSE=1
run cfg3a_coarse_b cfg3a SLC_ASM_COARSE=1
run cfg3a_direct cfg3a SLC_ASM_COARSE=0
run cfg3a_sort cfg3a SLC_ASM_BUCKET=0
run cfg3a_sort_2ms cfg3a SLC_ASM_BUCKET=0 SLC_BENCH_CLOCK_PERIOD=0.002
run cfg3a_trace cfg3a SLC_ASM_TRACE=1
tail -10 gpurun_out/r2_14_cfg3a_trace.err
run cfg3_mm cfg3
run cfg3_mm_2ms cfg3 SLC_BENCH_CLOCK_PERIOD=0.002
run cfg3t_bucket cfg3t SLC_TR_BUCKET=1
run cfg3t_sort cfg3t SLC_TR_BUCKET=0
run cfg5t_sort cfg5t SLC_TR_BUCKET=0
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_14_launches_cfg3a.csv python bench.py --workload cfg3a --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_14_ncu_cfg3a.log 2>&1
echo "ncu rc=$?"
python scripts/summarize_launches.py gpurun_out/r2_14_launches_cfg3a.csv | head -20
