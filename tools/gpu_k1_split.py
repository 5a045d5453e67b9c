"""K1 at the bench shape with and without the S_xx side kernel (how much of it hides behind the gradient units)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vgan_b200 import RBF, _lib
from vgan_b200.loss import mmd2_eval
N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
D = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
torch.manual_seed(0)
X = torch.randn(N, D, device="cuda"); U = torch.rand(N, D, device="cuda")
k = RBF(); bw, _ = k._bw_state(X.device)
for name, kw in (("grad+xx", {}), ("grad only", dict(xx_part=10**6 - 1, xx_nparts=10**6)), ("xx only (fwd)", dict(need_dY=False, row_end=128))):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ts = []
    for it in range(5):
        out = mmd2_eval(X, None, U, weight=0.0, bw_dev=bw, frozen=(k.bandwidth is not None), mults=k._mults,
                        precision="bf16", masked=True, k1_events=ev, **kw)
        torch.cuda.synchronize(); k._freeze()
        if it >= 2: ts.append(ev[0].elapsed_time(ev[1]))
    print(f"N={N} D={D} {name:14s}: K1 {sum(ts)/len(ts):8.3f} ms")
