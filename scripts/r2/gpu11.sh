#!/bin/bash
# round 2, GPU pass 11: where the 23 ms of the bucket assembly go (SLC_ASM_TRACE), early give-up flag,
# brick test expectation fixed, boundary SpMM rows spread over column tiles (2-GPU bench lines come later)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q --timeout 600 -p no:cacheprovider -k "brick or compress or transpose" > gpurun_out/r2_11_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_11_pytest.log
tail -6 gpurun_out/r2_11_pytest.log
run() {  # name, workload, env...
  name=$1; wl=$2; shift; shift
  env "$@" timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/r2_11_$name.json 2> gpurun_out/r2_11_$name.err
  echo "$name rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_11_$name.json'));print(d['ms_per_step'], d['roofline']['frac'], d.get('parity'))" 2>&1 | tail -1)"
}
run cfg3a_bucket cfg3a SLC_ASM_BUCKET=1
run cfg3a_trace cfg3a SLC_ASM_BUCKET=1 SLC_ASM_TRACE=1
tail -40 gpurun_out/r2_11_cfg3a_trace.err
run cfg4_bucket cfg4 SLC_ASM_BUCKET=1
run cfg4_sort cfg4 SLC_ASM_BUCKET=0
python - <<'PY'
import json
for n in ("bucket","sort"):
    d=json.load(open("gpurun_out/r2_11_cfg4_%s.json"%n)); print(n, d["ms_per_step"], d["config"].get("assembly"))
PY
