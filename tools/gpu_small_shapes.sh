# Launch list (ncu, per-kernel durations) of one process running three evaluations of each small / D <= 256 shape:
#   bash tools/gpu_small_shapes.sh [tag]   -> gpurun_out/<tag>_launches.csv ; read with tools/launch_summary.py
mkdir -p gpurun_out
TAG=${1:-r2n}
cd ${GRAFT_REPO_ROOT:-.}
SHAPES="4096 784 bf16 4096 49 fp32 4096 16 fp32 16384 128 bf16 16384 32 bf16 16384 128 fp32 65536 64 bf16 65536 256 bf16 500 20 fp32 16384 512 fp32"
timeout 120 python tools/gpu_shape_launches.py $SHAPES > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python tools/gpu_shape_launches.py $SHAPES > gpurun_out/${TAG}_ncu.log 2>&1
echo "launch list done: $TAG"
timeout 200 python tools/gpu_step_breakdown.py $SHAPES > gpurun_out/${TAG}_breakdown.jsonl 2> gpurun_out/${TAG}_breakdown.err; cat gpurun_out/${TAG}_breakdown.jsonl; tail -3 gpurun_out/${TAG}_breakdown.err
