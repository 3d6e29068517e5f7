set -x
cd $GRAFT_REPO_ROOT
N=4
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; head -c 250 gpurun_out/bench_${N}gpu.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/multigpu_probe.py ivf > gpurun_out/mg_ivf_${N}.json 2> gpurun_out/mg_ivf_${N}.err; grep "^{" gpurun_out/mg_ivf_${N}.json | cut -c1-400
