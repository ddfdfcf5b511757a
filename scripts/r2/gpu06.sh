#!/bin/bash
# round 2, GPU pass 6 (8 GPUs): 4-rank neighbour test, sharded handle over 8 devices, then the driver's
# scaling protocol N = 8, 4, 2, 1 (steps 20, warmup 5) and one 200-step run at N = 8
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2_06_gpus.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_multiproc.py tests/test_gpu_boundary.py -m gpu -q --timeout 500 -p no:cacheprovider -k "four_ranks or sharded_handle or second_device" > gpurun_out/r2_06_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_06_pytest.log
tail -4 gpurun_out/r2_06_pytest.log
for n in 8 4 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2_06_bench_n$n.json 2> gpurun_out/r2_06_bench_n$n.err
  echo "bench n$n rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_06_bench_n$n.json'));print(d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], [ (e['name'], e.get('ms_per_step')) for e in d.get('suite',[])])" 2>&1 | tail -1)"
done
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --suite none > gpurun_out/r2_06_bench_n1.json 2> gpurun_out/r2_06_bench_n1.err
echo "bench n1 rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_06_bench_n1.json'));print(d['ms_per_step'], d['value'])" 2>&1 | tail -1)"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29527 bench.py --gpus 8 --steps 200 --warmup 5 --suite none > gpurun_out/r2_06_bench_n8_200.json 2> gpurun_out/r2_06_bench_n8_200.err
echo "bench n8 200 steps rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_06_bench_n8_200.json'));print(d['ms_per_step'], d['value'])" 2>&1 | tail -1)"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29528 bench.py --gpus 8 --steps 20 --warmup 5 --suite none --no-graph > gpurun_out/r2_06_bench_n8_nograph.json 2> gpurun_out/r2_06_bench_n8_nograph.err
echo "bench n8 nograph rc=$? $(python -c "import json;d=json.load(open('gpurun_out/r2_06_bench_n8_nograph.json'));print(d['ms_per_step'], d['value'])" 2>&1 | tail -1)"
