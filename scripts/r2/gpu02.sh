#!/bin/bash
# round 2, GPU pass 2 (2 GPUs): multi-process / multi-device tests, then bench --gpus 2 with and without the graph
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2_02_gpus.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_multiproc.py tests/test_gpu_boundary.py tests/test_c_abi.py -m gpu -q --timeout 600 -p no:cacheprovider > gpurun_out/r2_02_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_02_pytest.log
tail -15 gpurun_out/r2_02_pytest.log
for extra in "" "--no-graph"; do
  tag=$(echo "$extra" | tr -d ' -')
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 5 $extra > gpurun_out/r2_02_bench_n2_$tag.json 2> gpurun_out/r2_02_bench_n2_$tag.err
  echo "bench n2 $extra rc=$?"
  tail -c 800 gpurun_out/r2_02_bench_n2_$tag.err
done
