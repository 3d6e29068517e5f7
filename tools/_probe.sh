set -x
cd $GRAFT_REPO_ROOT
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; tail -5 gpurun_out/pytest_gpu.log
P="python tools/scan_probe.py"
{

$P --rows 20000000 --dim 96 --batch 1024 --k 100 --dtype int8q --metric l2 --iters 6
$P --rows 100000000 --dim 96 --batch 1024 --k 100 --dtype int8q --metric l2 --iters 5
$P --rows 100000000 --dim 96 --batch 1024 --k 10 --dtype int8q --metric l2 --iters 5
$P --rows 100000000 --dim 96 --batch 128 --k 100 --dtype int8q --metric l2 --iters 5
$P --rows 100000000 --dim 96 --batch 32 --k 100 --dtype int8q --metric l2 --iters 5
$P --rows 100000000 --dim 96 --batch 8 --k 100 --dtype int8q --metric l2 --iters 5
$P --rows 100000000 --dim 96 --batch 1 --k 100 --dtype int8q --metric l2 --iters 8
$P --rows 20000000 --dim 96 --batch 1024 --k 100 --dtype fp16 --metric ip --iters 6
$P --rows 10000000 --dim 768 --batch 1024 --k 100 --dtype fp16 --metric ip --iters 6
$P --rows 1000000 --dim 128 --batch 1024 --k 10 --dtype fp32 --metric l2 --iters 8
$P --rows 1000000 --dim 128 --batch 10000 --k 10 --dtype fp32 --metric l2 --iters 8
} > gpurun_out/triage.log 2>&1
grep -v "^+" gpurun_out/triage.log | tail -40
