#!/bin/bash
# round 2, GPU pass 9 (8 GPUs): boundary blocks behind a lead of interior blocks, device-side rendezvous before
# the timed region; multi-process tests again; N = 8 lead sweep, then N = 8, 4, 2 with the default
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multiproc.py tests/test_gpu_sharded.py -m gpu -q --timeout 500 -p no:cacheprovider > gpurun_out/r2_09_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_09_pytest.log
tail -4 gpurun_out/r2_09_pytest.log
port=29600
run() {  # n, name, extra args..., env via SLC_*
  n=$1; name=$2; shift; shift
  port=$((port+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --warmup 5 "$@" > gpurun_out/r2_09_bench_$name.json 2> gpurun_out/r2_09_bench_$name.err
  echo "bench $name rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_09_bench_$name.json'));print(d['ms_per_step'], d['value'], d['roofline'].get('boundary_wait_us'), d['config']['timed_region'][-90:])" 2>&1 | tail -1)"
}
SLC_DIST_LEAD=0 run 8 n8_lead0 --steps 20 --suite none
SLC_DIST_LEAD=40 run 8 n8_lead40 --steps 20 --suite none
SLC_DIST_LEAD=70 run 8 n8_lead70 --steps 20 --suite none
SLC_DIST_LEAD=90 run 8 n8_lead90 --steps 20 --suite none
SLC_DIST_LEAD=40 run 8 n8_lead40_200 --steps 200 --suite none
run 4 n4 --steps 20 --suite none
run 2 n2 --steps 20 --suite none
run 8 n8_full --steps 20
