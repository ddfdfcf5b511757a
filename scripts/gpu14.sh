# round-1 pass 12 (N GPUs): sharded transpose / assembly over NCCL all-to-all-v, sharded GPU tests
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q > gpurun_out/p12_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/p12_pytest.log
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $T --master-port 29531 bench.py --gpus $N --workload cfg3t --steps 3 --warmup 1 > gpurun_out/p12_n${N}_cfg3t.json 2> gpurun_out/p12_n${N}_cfg3t.err
tail -c 1200 gpurun_out/p12_n${N}_cfg3t.json; tail -3 gpurun_out/p12_n${N}_cfg3t.err
timeout 600 $T --master-port 29532 bench.py --gpus $N --workload cfg3a --steps 3 --warmup 1 > gpurun_out/p12_n${N}_cfg3a.json 2> gpurun_out/p12_n${N}_cfg3a.err
tail -c 1200 gpurun_out/p12_n${N}_cfg3a.json; tail -3 gpurun_out/p12_n${N}_cfg3a.err
