#!/bin/bash
# round 2, GPU pass 10: bucket paths of compress and transpose (count / place / sort per column or row):
# the whole GPU suite (every plan is built through the transpose), then A/B bench lines and launch lists
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 -p no:cacheprovider > gpurun_out/r2_10_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_10_pytest.log
tail -12 gpurun_out/r2_10_pytest.log
run() {  # name, workload, env...
  name=$1; wl=$2; shift; shift
  env "$@" timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_10_$name.json 2> gpurun_out/r2_10_$name.err
  echo "$name rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_10_$name.json'));print(d['ms_per_step'], d['roofline']['frac'], d.get('parity'))" 2>&1 | tail -1)"
}
run cfg3a_bucket cfg3a SLC_ASM_BUCKET=1
run cfg3a_sort cfg3a SLC_ASM_BUCKET=0
run cfg3t_bucket cfg3t SLC_TR_BUCKET=1
run cfg3t_sort cfg3t SLC_TR_BUCKET=0
run cfg5t_bucket cfg5t SLC_TR_BUCKET=1
run cfg5t_sort cfg5t SLC_TR_BUCKET=0
run cfg4t_bucket cfg4t SLC_TR_BUCKET=1
run cfg4t_sort cfg4t SLC_TR_BUCKET=0
for w in cfg3a cfg5t; do
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_10_launches_$w.csv python bench.py --workload $w --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_10_ncu_$w.log 2>&1
  echo "ncu $w rc=$?"
done
for pdl in 1 0; do
  SLC_PDL=$pdl timeout 600 python bench.py --suite none --steps 20 --warmup 5 --no-cpu > gpurun_out/r2_10_pdl$pdl.json 2> gpurun_out/r2_10_pdl$pdl.err
  echo "pdl=$pdl rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_10_pdl$pdl.json'));print(d['ms_per_step'], d['e2e']['ms_per_step'])" 2>&1 | tail -1)"
done
