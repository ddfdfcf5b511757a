#!/bin/bash
# round 2, GPU pass 27 (2 GPUs): where the sharded transpose / assembly spend their time
mkdir -p gpurun_out
for w in cfg3t cfg3a; do
  SLC_SHARD_TRACE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 2 --workload $w --steps 5 --warmup 3 > gpurun_out/r2_27_$w.json 2> gpurun_out/r2_27_$w.err
  echo "$w rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_27_$w.json'));print(d['ms_per_step'], d['parity'])" 2>&1 | tail -1)"
  grep "slc sharded" gpurun_out/r2_27_$w.err | tail -9
done
for w in cfg3t cfg3a; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29572 bench.py --gpus 2 --workload $w --steps 5 --warmup 3 > gpurun_out/r2_27_${w}_plain.json 2> gpurun_out/r2_27_${w}_plain.err
  echo "$w plain rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_27_${w}_plain.json'));print(d['ms_per_step'], d['parity']['ok'])" 2>&1 | tail -1)"
done
