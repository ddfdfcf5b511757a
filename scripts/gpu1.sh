set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt; nproc >> gpurun_out/gpus.txt
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/b_cfg5.json 2> gpurun_out/b_cfg5.err; tail -c 2500 gpurun_out/b_cfg5.json
for w in cfg1 cfg2 cfg3 cfg5mm; do
  timeout 600 python bench.py --workload $w --no-cpu --steps 50 > gpurun_out/b_$w.json 2> gpurun_out/b_$w.err; tail -c 1500 gpurun_out/b_$w.json
done
( time timeout 900 python bench.py --workload cfg4 --no-cpu --steps 50 ) > gpurun_out/b_cfg4.json 2> gpurun_out/b_cfg4.err; tail -c 1500 gpurun_out/b_cfg4.json; tail -3 gpurun_out/b_cfg4.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/b_ref.json 2>&1; tail -c 800 gpurun_out/b_ref.json
for w in cfg5 cfg2 cfg3; do
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$w.csv python bench.py --workload $w --steps 3 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_$w.log 2>&1
done
ls -la gpurun_out
