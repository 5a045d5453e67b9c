import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vgan_b200 import MMDLossConstrained, RBF
from vgan_b200.loss import mmd2_eval
from oracle import mmd_oracle as orc

def run(N, D, prec, need_dX=False, grad=True, seed=0):
    rs = np.random.RandomState(seed)
    X = rs.randn(N, D).astype(np.float32); U = rs.rand(N, D).astype(np.float32)
    bw = orc.bandwidth_direct(np.vstack([X, U * X]))
    ref = orc.mmd2_terms(X, U * X, bw, need_dX=False)
    Xd, Ud = torch.tensor(X, device="cuda"), torch.tensor(U, device="cuda")
    k = RBF()
    bwd, frozen = k._bw_state(Xd.device)
    if need_dX:
        out = mmd2_eval(Xd, (Ud * Xd).contiguous(), None, weight=0.0, bw_dev=bwd, frozen=False, mults=k._mults,
                        precision=prec, need_dX=True, need_dY=True)
    else:
        out = mmd2_eval(Xd, None, Ud, weight=0.0, bw_dev=bwd, frozen=False, mults=k._mults, precision=prec,
                        need_dX=False, need_dY=grad, masked=True)
    torch.cuda.synchronize()
    s = out["sums"].cpu().numpy()
    print(f"N={N:5d} D={D:4d} {prec} dX={int(need_dX)} grad={int(grad)}: rel err sxx {(s[0]-ref[0])/ref[0]:+.2e} sxy {(s[1]-ref[1])/ref[1]:+.2e} syy {(s[2]-ref[2])/ref[2]:+.2e}")

for N in (256, 1024, 4096):
    for D in (64, 128):
        run(N, D, "bf16")
run(4096, 64, "fp32")
run(4096, 64, "bf16", need_dX=True)
run(4096, 64, "bf16", grad=False)
run(4096, 64, "bf16", seed=4160)
run(2048, 48, "bf16")
run(8192, 64, "bf16")
