#!/bin/bash
# round 2, GPU pass 18: coarse scatter held to 64 registers (4 blocks per SM)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --timeout 600 -p no:cacheprovider -k "compress" > gpurun_out/r2_18_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_18_pytest.log
tail -3 gpurun_out/r2_18_pytest.log
run() {  # name, workload, env...
  name=$1; wl=$2; shift; shift
  env "$@" timeout 600 python bench.py --workload $wl --steps 20 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_18_$name.json 2> gpurun_out/r2_18_$name.err
  echo "$name rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_18_$name.json'));print(d['ms_per_step'], d['roofline']['frac'], d.get('parity'))" 2>&1 | tail -1)"
}
run cfg3a cfg3a
run cfg3a_b cfg3a
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_18_launches_cfg3a.csv python bench.py --workload cfg3a --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_18_ncu_cfg3a.log 2>&1
python scripts/summarize_launches.py gpurun_out/r2_18_launches_cfg3a.csv | head -20
