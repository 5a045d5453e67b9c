# bench.py at 8 and 4 GPUs of one box (launched exactly as the driver does): bash tools/gpu_scale_run.sh [tag]
mkdir -p gpurun_out
TAG=${1:-r2n}
cd ${GRAFT_REPO_ROOT:-.}
for n in 8 4 2; do
  timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
    bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_n$n.json 2> gpurun_out/${TAG}_bench_n$n.err
  echo "n=$n rc=$?"; tail -c 1200 gpurun_out/${TAG}_bench_n$n.json; tail -c 400 gpurun_out/${TAG}_bench_n$n.err
done
