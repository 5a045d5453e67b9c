# r2n7: new fp32-class medium test, ncu full capture of the D <= 256 kernel, D_p = 256 on the pair kernel (variant)
mkdir -p gpurun_out; cd ${GRAFT_REPO_ROOT:-.}
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "medium or golden" 2>&1 | tail -3 | tee gpurun_out/r2n7_pytest.log
timeout 120 python tools/gpu_shape_launches.py 16384 128 bf16 > /dev/null 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:mmd2_tc_kernel -s 2 -c 1 -o gpurun_out/r2n7_tc_d128 -f python tools/gpu_shape_launches.py 16384 128 bf16 > gpurun_out/r2n7_ncu_full.log 2>&1
timeout 200 python tools/gpu_step_breakdown.py 65536 256 bf16 16384 256 bf16 4096 256 bf16 16384 200 bf16 > gpurun_out/r2n7_breakdown_prod.jsonl 2>&1; cat gpurun_out/r2n7_breakdown_prod.jsonl
cp vgan_b200/libvgan_b200_p256.so vgan_b200/libvgan_b200.so
timeout 200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "medium_sizes_bf16" 2>&1 | tail -3 | tee gpurun_out/r2n7_pytest_p256.log
timeout 200 python tools/gpu_step_breakdown.py 65536 256 bf16 16384 256 bf16 4096 256 bf16 16384 200 bf16 > gpurun_out/r2n7_breakdown_p256.jsonl 2>&1; cat gpurun_out/r2n7_breakdown_p256.jsonl
