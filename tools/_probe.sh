set -x
cd $GRAFT_REPO_ROOT
( time timeout 600 python -m pytest tests/test_sharding.py tests/test_ivf_gpu.py tests/test_plugin.py -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; tail -5 gpurun_out/pytest_gpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/multigpu_probe.py ivf > gpurun_out/mg_ivf_2.json 2> gpurun_out/mg_ivf_2.err; tail -c 700 gpurun_out/mg_ivf_2.json; tail -3 gpurun_out/mg_ivf_2.err
python tools/multigpu_probe.py ivf > gpurun_out/mg_ivf_1.json 2> gpurun_out/mg_ivf_1.err; tail -c 700 gpurun_out/mg_ivf_1.json
