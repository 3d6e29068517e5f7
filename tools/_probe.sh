set -x
cd $GRAFT_REPO_ROOT
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; tail -c 300 gpurun_out/bench_2gpu.json; tail -3 gpurun_out/bench_2gpu.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/multigpu_probe.py ivf > gpurun_out/mg_ivf_2.json 2> gpurun_out/mg_ivf_2.err; tail -c 600 gpurun_out/mg_ivf_2.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 tools/multigpu_probe.py mv > gpurun_out/mg_mv_2.json 2> gpurun_out/mg_mv_2.err; tail -c 600 gpurun_out/mg_mv_2.json
