set -x
cd $GRAFT_REPO_ROOT
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; tail -5 gpurun_out/pytest_gpu.log
python tools/ivf_probe.py 10000000 1536 16384 64 1,8,64,1024 > gpurun_out/ivf_c4.json 2> gpurun_out/ivf_c4.err; python -c "
import json; r=json.load(open('gpurun_out/ivf_c4.json'))
for x in r['results']: print(x['batch'], round(x['e2e_qps']), round(x['ms_per_call'],4), round(x['list_scan_ms'],4), round(x['list_scan_GBs']))"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"scan_|merge_|ivf_plan|prep_" -s 8 -c 8 --csv --log-file gpurun_out/ivf_launches.csv python tools/ivf_probe.py 10000000 1536 16384 64 1 > gpurun_out/ivf_ncu.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/ivf_launches.csv')) if len(r)>5]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value'); gi=hdr.index('Grid Size') if 'Grid Size' in hdr else None
for r in rows[1:25]: print(r[ki][:70], r[vi], r[gi] if gi is not None else '')
PY
P="python tools/scan_probe.py"
$P --rows 1000000 --dim 128 --batch 1 --k 10 --dtype fp32 --metric l2 --iters 10
$P --rows 10000000 --dim 768 --batch 1 --k 10 --dtype fp16 --metric l2 --iters 6
$P --rows 10000000 --dim 768 --batch 1 --k 10 --dtype fp16 --metric l2 --iters 6 --kernel 1
