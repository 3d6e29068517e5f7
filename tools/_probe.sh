set -x
cd $GRAFT_REPO_ROOT
python tools/multivector_probe.py > gpurun_out/mv_c5.json 2> gpurun_out/mv_c5.err; tail -c 300 gpurun_out/mv_c5.json
