"""K1 time of one rank's shard at the bench shape for world = 1, 2, 4, 8 (single GPU emulation of dist.py's split)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vgan_b200 import RBF
from vgan_b200.loss import mmd2_eval
N, D = 65536, 1024
torch.manual_seed(0)
X = torch.randn(N, D, device="cuda"); U = torch.rand(N, D, device="cuda")
k = RBF(); bw, _ = k._bw_state(X.device)
for world in (1, 2, 4, 8):
    for name, kw in (("grad+xx", dict(xx_part=0, xx_nparts=world)), ("grad only", dict(xx_part=10**6 - 1, xx_nparts=10**6))):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        ts = []
        for it in range(6):
            mmd2_eval(X, None, U, weight=0.0, bw_dev=bw, frozen=(k.bandwidth is not None), mults=k._mults,
                      precision="bf16", masked=True, k1_events=ev, row_begin=0, row_end=N // world, **kw)
            torch.cuda.synchronize(); k._freeze()
            if it >= 2: ts.append(ev[0].elapsed_time(ev[1]))
        t = sum(ts) / len(ts)
        print(f"world={world} {name:9s}: K1 {t:7.3f} ms  x world = {t * world:7.2f} ms")
# S_xx kernel alone for the same shards (forward only, no gradient units)
for world in (1, 8):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ts = []
    for it in range(5):
        mmd2_eval(X, None, U, weight=0.0, bw_dev=bw, frozen=True, mults=k._mults, precision="bf16", masked=True,
                  k1_events=ev, row_begin=0, row_end=N // world, xx_part=0, xx_nparts=world, need_dY=False)
        torch.cuda.synchronize()
        if it >= 2: ts.append(ev[0].elapsed_time(ev[1]))
    print(f"world={world} forward only (all sums, no gradient): {sum(ts)/len(ts):7.3f} ms")
