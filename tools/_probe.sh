set -x
cd $GRAFT_REPO_ROOT
( time timeout 900 python -m pytest tests/test_kmeans.py -x -q ) > gpurun_out/pytest_gpu.log 2>&1; tail -15 gpurun_out/pytest_gpu.log
