set -x
cd $GRAFT_REPO_ROOT
( time timeout 900 python -m pytest tests/test_flat_gpu.py tests/test_refine.py tests/test_kmeans.py tests/test_ivf_gpu.py -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; tail -4 gpurun_out/pytest_gpu.log
P="python tools/scan_probe.py"
{
$P --rows 20000000 --dim 96 --batch 1024 --k 100 --dtype fp16 --metric l2 --iters 6
$P --rows 20000000 --dim 96 --batch 1024 --k 10 --dtype fp16 --metric l2 --iters 6
ZVB200_MMA_DEBUG=128 $P --rows 20000000 --dim 96 --batch 1024 --k 10 --dtype fp16 --metric l2 --iters 3 | tail -3
$P --rows 10000000 --dim 768 --batch 1024 --k 10 --dtype fp16 --metric l2 --iters 6
$P --rows 1000000 --dim 128 --batch 1024 --k 10 --dtype fp32 --metric l2 --iters 8
$P --rows 1000000 --dim 128 --batch 10000 --k 10 --dtype fp32 --metric l2 --iters 8
$P --rows 10000000 --dim 128 --batch 1024 --k 10 --dtype fp32 --metric l2 --iters 8
$P --rows 1000000 --dim 1536 --batch 1024 --k 10 --dtype fp32 --metric cos --iters 8
python tools/kmeans_probe.py 1000000 1536 16384 2
} > gpurun_out/triage.log 2>&1
grep -v "^+" gpurun_out/triage.log | tail -16
