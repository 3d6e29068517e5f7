set -x
cd $GRAFT_REPO_ROOT
( timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; grep -E "passed|failed" gpurun_out/pytest_gpu.log
python tools/multivector_probe.py > gpurun_out/mv_c5.json 2> gpurun_out/mv_c5.err; tail -c 200 gpurun_out/mv_c5.json
