set -x
cd $GRAFT_REPO_ROOT
( time timeout 900 python -m pytest tests/test_flat_gpu.py -m gpu -x -q -k "qi8 or batch1 or small_batch or ragged" ) > gpurun_out/pytest_gpu.log 2>&1; tail -4 gpurun_out/pytest_gpu.log
P="python tools/scan_probe.py"
{
$P --rows 100000000 --dim 96 --batch 1 --k 100 --dtype int8q --metric l2 --iters 8
$P --rows 100000000 --dim 96 --batch 1 --k 10 --dtype int8q --metric l2 --iters 8
$P --rows 100000000 --dim 96 --batch 4 --k 100 --dtype int8q --metric l2 --iters 8
$P --rows 10000000 --dim 768 --batch 1 --k 10 --dtype int8q --metric l2 --iters 8
python tools/ivf_probe.py 10000000 1536 16384 64 1,8,64,1024
} > gpurun_out/triage.log 2>&1
grep -v "^+" gpurun_out/triage.log | tail -12
