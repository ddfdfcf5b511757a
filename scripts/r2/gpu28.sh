#!/bin/bash
# round 2, GPU pass 28 (2 GPUs): long runs of empty buckets filled in parallel; flat scan of the piece pointers
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py tests/test_gpu_multiproc.py tests/test_gpu_boundary.py -m gpu -x -q --timeout 300 -p no:cacheprovider -k "transpose or ctrans or sharded or stacking or mulv" > gpurun_out/r2_28_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_28_pytest.log
tail -4 gpurun_out/r2_28_pytest.log
for w in cfg3t; do
  SLC_SHARD_TRACE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 2 --workload $w --steps 5 --warmup 3 > gpurun_out/r2_28_$w.json 2> gpurun_out/r2_28_$w.err
  echo "$w rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_28_$w.json'));print(d['ms_per_step'], d['parity']['ok'])" 2>&1 | tail -1)"
  grep "slc sharded" gpurun_out/r2_28_$w.err | tail -7
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29572 bench.py --gpus 2 --workload $w --steps 5 --warmup 3 > gpurun_out/r2_28_${w}_plain.json 2> gpurun_out/r2_28_${w}_plain.err
  echo "$w plain rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_28_${w}_plain.json'));print(d['ms_per_step'], d['parity']['ok'])" 2>&1 | tail -1)"
done
timeout 300 python bench.py --workload cfg5t --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2_28_cfg5t.json 2>/dev/null; python -c "import json;d=json.load(open('gpurun_out/r2_28_cfg5t.json'));print('cfg5t', d['ms_per_step'])"
