set -x
cd $GRAFT_REPO_ROOT
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; tail -4 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -7 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; head -c 300 gpurun_out/bench_1gpu.json; wc -l gpurun_out/bench_1gpu.json
ncu --set full --clock-control none --import-source on -k regex:scan_mma -s 2 -c 1 -f -o gpurun_out/prof_qi8_shadow python tools/scan_probe.py --rows 20000000 --dim 96 --batch 1024 --k 100 --dtype int8q --metric l2 --iters 4 > gpurun_out/ncu_shadow.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:scan_mma -s 2 -c 1 -f -o gpurun_out/prof_tf32 python tools/scan_probe.py --rows 1000000 --dim 128 --batch 1024 --k 10 --dtype fp32 --metric l2 --iters 4 > gpurun_out/ncu_tf32.log 2>&1
