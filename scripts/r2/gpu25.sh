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
run cfg2_single cfg2 SLC_SPMM_CLUSTER=0
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --timeout 300 -p no:cacheprovider -k "window" > gpurun_out/r2_25_pytest.log 2>&1
tail -3 gpurun_out/r2_25_pytest.log
