#!/bin/bash
# round 2, GPU pass 7: planes SpMM block shapes + ncu at the default; SpGEMM input staging variants + ncu;
# launch lists of the transposes / assembly
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 600 -p no:cacheprovider -x -k "brick or mm_" > gpurun_out/r2_07_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_07_pytest.log
tail -5 gpurun_out/r2_07_pytest.log
run() {  # workload, name, env...
  wl=$1; name=$2; shift; shift
  env "$@" timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 --no-e2e > gpurun_out/r2_07_${wl}_$name.json 2> gpurun_out/r2_07_${wl}_$name.err
  echo "$wl $name rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_07_${wl}_$name.json'));print(d['ms_per_step'], d['roofline']['frac'], d['parity']['ok'])" 2>&1 | tail -1)"
}
run cfg5mm r512j1 SLC_PLANES_ROWS=512
run cfg5mm r256j1 SLC_PLANES_ROWS=256
run cfg5mm r256j2 SLC_PLANES_ROWS=256 SLC_PLANES_JT=2
run cfg5mm r512j1s4 SLC_PLANES_ROWS=512 SLC_PLANES_STAGES=4
run cfg5mm r512j1s6 SLC_PLANES_ROWS=512 SLC_PLANES_STAGES=6
run cfg3 a256t64 SLC_MM_ACAP=256
run cfg3 a256t128 SLC_MM_ACAP=256 SLC_MM_THREADS=128
run cfg3 a256t32 SLC_MM_ACAP=256 SLC_MM_THREADS=32
run cfg3 a0 SLC_MM_ACAP=0
run cfg3 a512t64 SLC_MM_ACAP=512
ncu --set full --clock-control none --import-source on -k regex:k_spmm_planes -c 1 -o gpurun_out/r2_07_ncu_planes python bench.py --workload cfg5mm --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_07_ncu_planes.log 2>&1
echo "ncu planes rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_mm_merge -c 4 -o gpurun_out/r2_07_ncu_mm python bench.py --workload cfg3 --steps 2 --warmup 1 --no-e2e --no-cpu > gpurun_out/r2_07_ncu_mm.log 2>&1
echo "ncu mm rc=$?"
for wl in cfg5t cfg4t cfg3a; do
  ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_07_launches_$wl.csv python bench.py --workload $wl --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_07_launches_$wl.log 2>&1
  echo "launches $wl rc=$?"
done
