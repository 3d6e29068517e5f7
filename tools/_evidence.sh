mkdir -p gpurun_out
python tools/scan_probe.py --rows 100000000 --dim 96 --batch 1 --k 100 --dtype int8q --metric l2 --iters 8 2>&1 | tail -1
ncu --set full --clock-control none --import-source on -k regex:scan_gemv -s 2 -c 1 -o gpurun_out/r02_c3_b1_gemv_after python tools/scan_probe.py --rows 100000000 --dim 96 --batch 1 --k 100 --dtype int8q --metric l2 --iters 3 > gpurun_out/c3_b1_after_ncu.log 2>&1
tail -1 gpurun_out/c3_b1_after_ncu.log
