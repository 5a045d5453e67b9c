mkdir -p gpurun_out; cd ${GRAFT_REPO_ROOT:-.}
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/r2n12_pytest.log
timeout 200 python tools/gpu_step_breakdown.py 4096 784 bf16 16384 128 bf16 16384 32 bf16 65536 64 bf16 4096 49 fp32 4096 16 fp32 500 20 fp32 4096 256 bf16 2048 784 bf16 > gpurun_out/r2n12_breakdown.jsonl 2>&1; cat gpurun_out/r2n12_breakdown.jsonl
timeout 200 python tools/gpu_cluster_timers.py 4096 784 --no-build > gpurun_out/r2n12_timers_4096_784.txt 2>&1; cat gpurun_out/r2n12_timers_4096_784.txt | tail -9
timeout 200 python tools/gpu_cluster_timers.py 65536 1024 --no-build > gpurun_out/r2n12_timers_65536_1024.txt 2>&1; cat gpurun_out/r2n12_timers_65536_1024.txt | tail -9
timeout 400 python bench.py --steps 8 --warmup 3 --no-secondary > gpurun_out/r2n12_bench.json 2> gpurun_out/r2n12_bench.err; cut -c1-400 gpurun_out/r2n12_bench.json
