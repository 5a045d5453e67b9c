"""K1 time (CUDA events around vgan_b200_mmd2_fwd_bwd) of the D_p <= 256 kernel at epilogue-bound shapes."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vgan_b200 import RBF
from vgan_b200.loss import mmd2_eval
for N, D in ((16384, 32), (16384, 128), (16384, 256), (65536, 64), (65536, 256), (131072, 128)):
    torch.manual_seed(0)
    X = torch.randn(N, D, device="cuda"); U = torch.rand(N, D, device="cuda")
    k = RBF(); bw, _ = k._bw_state(X.device)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ts = []
    for it in range(8):
        mmd2_eval(X, None, U, weight=0.0, bw_dev=bw, frozen=(k.bandwidth is not None), mults=k._mults,
                  precision="bf16", masked=True, k1_events=ev)
        torch.cuda.synchronize(); k._freeze()
        if it >= 3: ts.append(ev[0].elapsed_time(ev[1]))
    t = sum(ts) / len(ts)
    print(f"N={N:6d} D={D:4d}: K1 {t:8.3f} ms   {2.5 * N * N / t / 1e9:7.2f} G pairs/ms... {2.5*N*N/(t*1e-3)/148/1.9e9:5.2f} pairs/clk/SM @1.9GHz", flush=True)
