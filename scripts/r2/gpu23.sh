#!/bin/bash
# round 2, GPU pass 23: cluster window SpMM (two CTAs share the union of their windows through TMA multicast)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_boundary.py -m gpu -x -q --timeout 300 -p no:cacheprovider -k "window or feast or mulm" > gpurun_out/r2_23_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_23_pytest.log
tail -8 gpurun_out/r2_23_pytest.log
run() {  # name, workload, env...
  name=$1; wl=$2; shift; shift
  env "$@" timeout 300 python bench.py --workload $wl --steps 20 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_23_$name.json 2> gpurun_out/r2_23_$name.err
  echo "$name rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_23_$name.json'));print(d['ms_per_step'], d['roofline']['frac'], d.get('parity'))" 2>&1 | tail -1)"
}
run cfg2_cluster cfg2 SLC_SPMM_CLUSTER=1
run cfg2_single cfg2 SLC_SPMM_CLUSTER=0
run cfg2_cluster_b cfg2 SLC_SPMM_CLUSTER=1
