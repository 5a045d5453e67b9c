"""Start/end (globaltimer) of the gradient kernel and the S_xx kernel inside one K1 call: do they overlap?"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vgan_b200 import RBF, _lib
from vgan_b200.loss import mmd2_eval
N, D = 65536, 1024
world = int(sys.argv[1]) if len(sys.argv) > 1 else 1
torch.manual_seed(0)
X = torch.randn(N, D, device="cuda"); U = torch.rand(N, D, device="cuda")
k = RBF(); bw, _ = k._bw_state(X.device)
L = _lib.lib()
tl = torch.zeros(4, dtype=torch.int64, device="cuda")
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
for it in range(6):
    tl.copy_(torch.tensor([2**62, 0, 2**62, 0]))
    L.vgan_b200_debug_timeline(tl.data_ptr())
    mmd2_eval(X, None, U, weight=0.0, bw_dev=bw, frozen=(k.bandwidth is not None), mults=k._mults, precision="bf16",
              masked=True, k1_events=ev, row_begin=0, row_end=N // world, xx_part=0, xx_nparts=world)
    torch.cuda.synchronize(); k._freeze()
    t = tl.cpu().numpy().astype(float); t0 = min(t[0], t[2])
    print(f"  it {it}: K1 events {ev[0].elapsed_time(ev[1]):7.3f} ms | gradient {1e-6*(t[0]-t0):6.3f}..{1e-6*(t[1]-t0):6.3f}  S_xx {1e-6*(t[2]-t0):6.3f}..{1e-6*(t[3]-t0):6.3f}")
L.vgan_b200_debug_timeline(None)
t = tl.cpu().numpy().astype(float)
t0 = min(t[0], t[2])
print(f"world={world}: gradient kernel {1e-6*(t[0]-t0):7.3f} .. {1e-6*(t[1]-t0):7.3f} ms ; S_xx kernel {1e-6*(t[2]-t0):7.3f} .. {1e-6*(t[3]-t0):7.3f} ms")
