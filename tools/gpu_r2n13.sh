mkdir -p gpurun_out; cd ${GRAFT_REPO_ROOT:-.}
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -12 | tee gpurun_out/r2n13_pytest.log
timeout 200 python tools/gpu_step_breakdown.py 16384 32 bf16 16384 100 bf16 16384 128 bf16 16384 200 bf16 65536 60 bf16 4096 49 fp32 4096 16 fp32 500 20 fp32 > gpurun_out/r2n13_breakdown.jsonl 2>&1; cat gpurun_out/r2n13_breakdown.jsonl
