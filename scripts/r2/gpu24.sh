#!/bin/bash
# round 2, GPU pass 25: cluster window kernel with relaxed remote arrives
mkdir -p gpurun_out
run() {  # name, workload, env...
  name=$1; wl=$2; shift; shift
  env "$@" timeout 300 python bench.py --workload $wl --steps 20 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_25_$name.json 2> gpurun_out/r2_25_$name.err
  echo "$name rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_25_$name.json'));print(d['ms_per_step'], d['roofline']['frac'], d.get('parity'))" 2>&1 | tail -1)"
}
run cfg2_mode1 cfg2 SLC_SPMM_CLUSTER=1 SLC_WIN2_MODE=1
run cfg2_mode2 cfg2 SLC_SPMM_CLUSTER=1 SLC_WIN2_MODE=2
run cfg2_mode0 cfg2 SLC_SPMM_CLUSTER=1 SLC_WIN2_MODE=0
timeout 300 ncu --set full --clock-control none -k regex:k_spmm_window2 -s 2 -c 1 -o gpurun_out/r2_25_ncu_window2 python bench.py --workload cfg2 --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_25_ncu.log 2>&1
echo "ncu rc=$?"
