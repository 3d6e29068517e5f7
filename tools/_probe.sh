# scratch: the command list of the latest gpurun call (see profiles/ for what was kept)
set -x
cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
