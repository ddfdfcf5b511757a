#!/bin/bash
# round 2, GPU pass 13: host-side timeline of the coarse assembly path (no synchronisation between phases)
mkdir -p gpurun_out
run() {  # name, workload, env...
  name=$1; wl=$2; shift; shift
  env "$@" timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_13_$name.json 2> gpurun_out/r2_13_$name.err
  echo "$name rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_13_$name.json'));print(d['ms_per_step'], d['roofline']['frac'], d.get('parity'))" 2>&1 | tail -1)"
}
run coarse_host cfg3a SLC_ASM_COARSE=1 SLC_ASM_TRACE=2
tail -26 gpurun_out/r2_13_coarse_host.err
run coarse_blocking cfg3a SLC_ASM_COARSE=1 CUDA_LAUNCH_BLOCKING=1
run coarse_plain cfg3a SLC_ASM_COARSE=1
run sort_host cfg3a SLC_ASM_BUCKET=0
