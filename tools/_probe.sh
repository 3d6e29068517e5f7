set -x
cd $GRAFT_REPO_ROOT
timeout 300 python -m pytest tests/test_kmeans.py -x -q 2>&1 | tail -8
