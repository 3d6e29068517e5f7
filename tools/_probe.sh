set -x
cd $GRAFT_REPO_ROOT
( time timeout 900 python -m pytest tests/test_flat_gpu.py tests/test_refine.py -m gpu -x -q -k "qi8 or refine or determin" ) > gpurun_out/pytest_gpu.log 2>&1; tail -4 gpurun_out/pytest_gpu.log
P="python tools/scan_probe.py"
{
ZVB200_MMA_DEBUG=128 $P --rows 20000000 --dim 96 --batch 1024 --k 100 --dtype int8q --metric l2 --iters 3 | tail -3
$P --rows 20000000 --dim 96 --batch 1024 --k 100 --dtype int8q --metric l2 --iters 6
$P --rows 20000000 --dim 96 --batch 1024 --k 10 --dtype int8q --metric l2 --iters 6
$P --rows 100000000 --dim 96 --batch 1024 --k 100 --dtype int8q --metric l2 --iters 5
$P --rows 100000000 --dim 96 --batch 1024 --k 10 --dtype int8q --metric l2 --iters 5
$P --rows 2000000 --dim 96 --batch 1024 --k 100 --dtype int8q --metric l2 --iters 8
} > gpurun_out/triage.log 2>&1
grep -v "^+" gpurun_out/triage.log | tail -16
