set -x
cd $GRAFT_REPO_ROOT
python bench.py > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; tail -c 600 gpurun_out/bench_1gpu.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; tail -c 400 gpurun_out/bench_ref.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 > gpurun_out/ncu_bench.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:scan_mma -s 2 -c 1 -f -o gpurun_out/prof_qi8_shadow python tools/scan_probe.py --rows 20000000 --dim 96 --batch 1024 --k 100 --dtype int8q --metric l2 --iters 4 > gpurun_out/ncu_shadow.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:scan_mma -s 2 -c 1 -f -o gpurun_out/prof_f16_b1024 python tools/mma_probe.py 10000000 1024 768 10 > gpurun_out/ncu_f16.log 2>&1
ls -la gpurun_out/*.ncu-rep
