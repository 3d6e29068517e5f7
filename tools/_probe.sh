set -x
cd $GRAFT_REPO_ROOT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; tail -5 gpurun_out/pytest_gpu.log
P="python tools/scan_probe.py"
{
for k in 10 100; do
  ZVB200_MMA_DEBUG=128 $P --rows 20000000 --dim 96 --batch 1024 --k $k --dtype int8q --metric l2 --iters 4
  $P --rows 20000000 --dim 96 --batch 1024 --k $k --dtype int8q --metric l2 --iters 6
  for l in 0 1 2; do ZVB200_LPR_LOG2=$l $P --rows 20000000 --dim 96 --batch 1 --k $k --dtype int8q --metric l2 --iters 8; done
done
ZVB200_MMA_DEBUG=1 $P --rows 20000000 --dim 96 --batch 1024 --k 10 --dtype int8q --metric l2 --iters 6
ZVB200_MMA_DEBUG=16 $P --rows 20000000 --dim 96 --batch 1024 --k 100 --dtype int8q --metric l2 --iters 6
$P --rows 20000000 --dim 96 --batch 1024 --k 100 --dtype int8 --metric ip --iters 6
$P --rows 20000000 --dim 96 --batch 1024 --k 100 --dtype fp16 --metric ip --iters 6
$P --rows 20000000 --dim 96 --batch 1024 --k 100 --dtype fp16 --metric l2 --iters 6
$P --rows 20000000 --dim 96 --batch 1024 --k 10 --dtype fp16 --metric ip --iters 6
ZVB200_MMA_DEBUG=128 $P --rows 20000000 --dim 96 --batch 1024 --k 100 --dtype fp16 --metric ip --iters 4
} > gpurun_out/triage.log 2>&1
grep -v "^+" gpurun_out/triage.log | tail -60
