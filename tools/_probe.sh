set -x
cd $GRAFT_REPO_ROOT
for i in 1 2 3 4; do timeout 300 python -m pytest tests/test_flat_gpu.py tests/test_plugin.py -m gpu -x -q -k "concurrent or many_threads or race" 2>&1 | tail -1; done
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; grep -E "passed|failed" gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; wc -l gpurun_out/bench_1gpu.json; head -c 200 gpurun_out/bench_1gpu.json
