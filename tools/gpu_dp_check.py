"""Two (or more) ranks on one box: dist.data_parallel_masked_step and dist.sharded_masked_step against the single-rank
evaluation of the same global batch, at widths served by the D <= 256 kernel (ones columns, column-range cuts) and by
the pair kernel.  Launch: python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/gpu_dp_check.py"""
import os, sys
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vgan_b200 import MMDLossConstrained, RBF
from vgan_b200.dist import data_parallel_masked_step, sharded_masked_step

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl")
ok = True
for N, D in ((4096, 96), (2048, 200), (4096, 64), (2048, 320)):
    g = torch.Generator(device="cuda").manual_seed(7)
    X = torch.randn(N, D, device="cuda", generator=g)
    U = torch.rand(N, D, device="cuda", generator=g)
    ref = sharded_masked_step(MMDLossConstrained(10.0, kernel=RBF(), precision="bf16"), X, U, 0, 1)     # whole batch, this GPU
    n = N // world
    b, e = rank * n, (rank + 1) * n
    out = data_parallel_masked_step(MMDLossConstrained(10.0, kernel=RBF(), precision="bf16"), X[b:e].clone(), U[b:e].clone())
    err_dp = float((out["dU"] - ref["dU"][b:e]).norm() / ref["dU"][b:e].norm())
    out2 = sharded_masked_step(MMDLossConstrained(10.0, kernel=RBF(), precision="bf16"), X, U, rank, world)
    rb, re_ = out2["rows"]
    err_sh = float((out2["dU"] - ref["dU"][rb:re_]).norm() / ref["dU"][rb:re_].norm())
    dl = abs(float(out["loss"]) - float(ref["loss"])), abs(float(out2["loss"]) - float(ref["loss"]))
    cs = float((out["dU_colsum"] - ref["dU_colsum"]).abs().max() / ref["dU_colsum"].abs().max())
    good = err_dp <= 1e-5 and err_sh <= 1e-5 and max(dl) <= 1e-6 and cs <= 1e-5
    ok = ok and good
    print(f"rank {rank} N={N} D={D}: dU err data-parallel {err_dp:.1e} sharded {err_sh:.1e}  loss diff {dl[0]:.1e} {dl[1]:.1e}  colsum {cs:.1e}  {'OK' if good else 'MISMATCH'}", flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
