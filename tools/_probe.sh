set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/scan_probe.py --rows 10000000 --dim 768 --batch 1024 --dtype int8 --metric l2 --iters 5 2>&1 | tail -1
python tools/scan_probe.py --rows 10000000 --dim 768 --batch 1024 --dtype int8 --metric ip --iters 5 2>&1 | tail -1
python tools/scan_probe.py --rows 10000000 --dim 768 --batch 1024 --dtype int8q --metric ip --iters 5 2>&1 | tail -1
