set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
N=${1:-2}
for sc in strong weak; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 200 --warmup 10 --scaling $sc > gpurun_out/b_n${N}_$sc.json 2> gpurun_out/b_n${N}_$sc.err
tail -c 2500 gpurun_out/b_n${N}_$sc.json; tail -5 gpurun_out/b_n${N}_$sc.err
done
SLC_BARRIER=nccl timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 200 --warmup 10 --no-e2e > gpurun_out/b_n${N}_ncclbar.json 2> gpurun_out/b_n${N}_ncclbar.err
tail -c 1500 gpurun_out/b_n${N}_ncclbar.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/b_n${N}_ref.json 2>&1
tail -c 600 gpurun_out/b_n${N}_ref.json
