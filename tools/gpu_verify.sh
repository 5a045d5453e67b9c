# Verification run of a round state on one B200: GPU tests, smoke, default bench line (with `secondary`), the CPU reference
# arm, the launch list of the bench command and ncu --set full captures of the two gradient kernels.
#   bash tools/gpu_verify.sh <tag>   -> gpurun_out/<tag>_*
mkdir -p gpurun_out; cd ${GRAFT_REPO_ROOT:-.}
TAG=${1:-verify}
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/${TAG}_pytest.log; tail -2 gpurun_out/${TAG}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -4 gpurun_out/${TAG}_smoke.log
timeout 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; cut -c1-300 gpurun_out/${TAG}_bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; cut -c1-300 gpurun_out/${TAG}_bench_ref.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-secondary > gpurun_out/${TAG}_ncu_launches.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:mmd2_tc_pair_kernel -s 3 -c 1 -f -o gpurun_out/${TAG}_pair_k1 python bench.py --steps 2 --warmup 3 --no-secondary > gpurun_out/${TAG}_ncu_pair.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:mmd2_tc_kernel -s 2 -c 1 -f -o gpurun_out/${TAG}_tc_d32 python tools/gpu_shape_launches.py 16384 32 bf16 > gpurun_out/${TAG}_ncu_tc.log 2>&1
ls -la gpurun_out | grep ${TAG}
