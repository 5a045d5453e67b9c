# Short verification of a state on one B200: GPU tests, smoke, default bench line.  bash tools/gpu_verify_short.sh <tag>
mkdir -p gpurun_out; cd ${GRAFT_REPO_ROOT:-.}
TAG=${1:-verify}
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/${TAG}_pytest.log; tail -1 gpurun_out/${TAG}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -4 gpurun_out/${TAG}_smoke.log
timeout 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; cut -c1-220 gpurun_out/${TAG}_bench.json
