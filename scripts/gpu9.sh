# round-1 pass 7 (N GPUs): row-partitioned SpGEMM, all-gather SpMV on R-MAT, stencil SpMV/SpMM
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
N=${1:-2}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $T --master-port 29521 bench.py --gpus $N --workload cfg3 --steps 10 --warmup 3 > gpurun_out/p7_n${N}_cfg3.json 2> gpurun_out/p7_n${N}_cfg3.err
tail -c 1500 gpurun_out/p7_n${N}_cfg3.json; tail -3 gpurun_out/p7_n${N}_cfg3.err
timeout 900 $T --master-port 29522 bench.py --gpus $N --workload cfg4 --grid ${2:-22} --steps 50 --warmup 5 > gpurun_out/p7_n${N}_cfg4.json 2> gpurun_out/p7_n${N}_cfg4.err
tail -c 1800 gpurun_out/p7_n${N}_cfg4.json; tail -3 gpurun_out/p7_n${N}_cfg4.err
timeout 600 $T --master-port 29523 bench.py --gpus $N --steps 200 --warmup 10 > gpurun_out/p7_n${N}_spmv.json 2> gpurun_out/p7_n${N}_spmv.err
tail -c 1200 gpurun_out/p7_n${N}_spmv.json; tail -3 gpurun_out/p7_n${N}_spmv.err
