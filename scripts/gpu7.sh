# round-1 pass 5 (N GPUs): row-partitioned SpMV after the refactor, row-partitioned SpMM
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
N=${1:-2}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $T --master-port 29511 bench.py --gpus $N --steps 200 --warmup 10 > gpurun_out/b_n${N}_spmv.json 2> gpurun_out/b_n${N}_spmv.err
tail -c 1800 gpurun_out/b_n${N}_spmv.json; tail -3 gpurun_out/b_n${N}_spmv.err
timeout 900 $T --master-port 29512 bench.py --gpus $N --workload cfg5mm --steps 20 --warmup 3 > gpurun_out/b_n${N}_spmm.json 2> gpurun_out/b_n${N}_spmm.err
tail -c 1800 gpurun_out/b_n${N}_spmm.json; tail -5 gpurun_out/b_n${N}_spmm.err
timeout 900 $T --master-port 29513 bench.py --gpus $N --workload cfg5mm --scaling weak --steps 20 --warmup 3 --no-e2e > gpurun_out/b_n${N}_spmm_weak.json 2> gpurun_out/b_n${N}_spmm_weak.err
tail -c 1500 gpurun_out/b_n${N}_spmm_weak.json; tail -5 gpurun_out/b_n${N}_spmm_weak.err
