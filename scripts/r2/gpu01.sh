#!/bin/bash
# round 2, GPU pass 1: all GPU tests (new boundary / multiproc / config tests included), then the default bench
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2_01_gpus.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -p no:cacheprovider > gpurun_out/r2_01_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_01_pytest.log
tail -30 gpurun_out/r2_01_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_01_bench.json 2> gpurun_out/r2_01_bench.err
echo "bench rc=$?"
tail -c 1500 gpurun_out/r2_01_bench.err
