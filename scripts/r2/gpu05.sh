#!/bin/bash
# round 2, GPU pass 5: TMA-staged planes SpMM (tests, variants, ncu); transpose with fused pointers / tiled
# expand; e2e pipeline tapered at both ends
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py tests/test_gpu_properties.py -m gpu -q --timeout 600 -p no:cacheprovider -x -k "brick or block_product or transpose or ctrans or hermitian or mulm or spmm" > gpurun_out/r2_05_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_05_pytest.log
tail -5 gpurun_out/r2_05_pytest.log
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 600 python bench.py --workload cfg5mm --steps 10 --warmup 3 --no-e2e > gpurun_out/r2_05_mm_$name.json 2> gpurun_out/r2_05_mm_$name.err
  echo "$name rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_05_mm_$name.json'));print(d['ms_per_step'], d['roofline']['frac'], d['parity']['ok'], d['config']['plan'][-8:])" 2>&1 | tail -1)"
}
run j2 SLC_PLANES_JT=2
run j1 SLC_PLANES_JT=1
run j4 SLC_PLANES_JT=4
run j2s2 SLC_PLANES_JT=2 SLC_PLANES_STAGES=2
run j1s4 SLC_PLANES_JT=1 SLC_PLANES_STAGES=4
run panel SLC_SPMM_BRICK=0
ncu --set full --clock-control none --import-source on -k regex:k_spmm_planes -c 1 -o gpurun_out/r2_05_ncu_planes python bench.py --workload cfg5mm --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_05_ncu.log 2>&1
echo "ncu rc=$?"
for wl in cfg5t cfg3t cfg3a cfg4t; do
  timeout 900 python bench.py --workload $wl --steps 10 --warmup 3 --no-e2e > gpurun_out/r2_05_$wl.json 2> gpurun_out/r2_05_$wl.err
  echo "$wl rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_05_$wl.json'));print(d['ms_per_step'], d['roofline']['frac'], d['parity']['ok'])" 2>&1 | tail -1)"
done
e2e() {
  name=$1; shift
  env "$@" timeout 600 python bench.py --suite none --steps 20 --warmup 5 --no-cpu > gpurun_out/r2_05_e2e_$name.json 2> gpurun_out/r2_05_e2e_$name.err
  echo "e2e $name rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_05_e2e_$name.json'));print(d['ms_per_step'], d['e2e']['ms_per_step'])" 2>&1 | tail -1)"
}
e2e both8 SLC_E2E_TAPER=2
e2e both6 SLC_E2E_TAPER=2 SLC_E2E_CHUNKS=6
e2e both12 SLC_E2E_TAPER=2 SLC_E2E_CHUNKS=12
e2e end8 SLC_E2E_TAPER=1
e2e both8m1 SLC_E2E_TAPER=2 SLC_E2E_MODE=1
e2e trace SLC_E2E_TRACE=1
tail -14 gpurun_out/r2_05_e2e_trace.err
