#!/bin/bash
# round 2, GPU pass 8: two-rows-per-thread planes SpMM (tests, variants, ncu), sharded SpMM test
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py tests/test_gpu_configs.py -m gpu -q --timeout 600 -p no:cacheprovider -x -k "brick or block_product or spmm or mulm or cfg5" > gpurun_out/r2_08_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_08_pytest.log
tail -5 gpurun_out/r2_08_pytest.log
run() {  # workload, name, env...
  wl=$1; name=$2; shift; shift
  env "$@" timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 --no-e2e > gpurun_out/r2_08_${wl}_$name.json 2> gpurun_out/r2_08_${wl}_$name.err
  echo "$wl $name rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_08_${wl}_$name.json'));print(d['ms_per_step'], d['roofline']['frac'], d['parity']['ok'])" 2>&1 | tail -1)"
}
run cfg5mm pair384 SLC_PLANES_PAIR=1
run cfg5mm pair256 SLC_PLANES_PAIR=1 SLC_PLANES_CONS=256
run cfg5mm pair384s4 SLC_PLANES_PAIR=1 SLC_PLANES_STAGES=4
run cfg5mm one512 SLC_PLANES_PAIR=0
ncu --set full --clock-control none --import-source on -k regex:k_spmm_planes -c 1 -o gpurun_out/r2_08_ncu_planes2 python bench.py --workload cfg5mm --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_08_ncu_planes2.log 2>&1
echo "ncu planes2 rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2_08_launches_cfg5mm.csv python bench.py --workload cfg5mm --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_08_launches_cfg5mm.log 2>&1
echo "launches rc=$?"
