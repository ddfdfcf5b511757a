#!/bin/bash
# round 2, GPU pass 21 (2 GPUs): halo staging for the boundary rows of the row-partitioned block product
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_multiproc.py tests/test_gpu_boundary.py -m gpu -x -q --timeout 600 -p no:cacheprovider > gpurun_out/r2_21_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_21_pytest.log
tail -5 gpurun_out/r2_21_pytest.log
for halo in 1 0; do
  SLC_SPMM_HALO=$halo timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2955$halo bench.py --gpus 2 --workload cfg5mm --steps 20 --warmup 5 > gpurun_out/r2_21_mm_halo$halo.json 2> gpurun_out/r2_21_mm_halo$halo.err
  echo "halo=$halo rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_21_mm_halo$halo.json'));print(d['ms_per_step'], d['parity'])" 2>&1 | tail -1)"
done
