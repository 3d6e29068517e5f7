set -x
python -m pytest tests/test_refine.py -m gpu -x -q 2>&1 | tail -25
