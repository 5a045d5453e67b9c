mkdir -p gpurun_out; cd ${GRAFT_REPO_ROOT:-.}
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/r2q_pytest.log
timeout 200 python tools/gpu_step_breakdown.py 65536 64 bf16 16384 64 bf16 4096 64 bf16 16384 63 bf16 16384 60 bf16 16384 128 bf16 > gpurun_out/r2q_breakdown.jsonl 2>&1; cat gpurun_out/r2q_breakdown.jsonl
