"""Per-role wait timers of the pair kernel (cycles spent waiting on each barrier).  Needs a library built with
`make -C vgan_b200/csrc clean all TIMERS=1` — the product build compiles the clock reads out."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vgan_b200 import RBF, _lib
from vgan_b200.loss import mmd2_eval
N, D = int(sys.argv[1]) if len(sys.argv) > 1 else 16384, 1024
torch.manual_seed(0)
X = torch.randn(N, D, device="cuda"); U = torch.rand(N, D, device="cuda")
k = RBF(); bw, _ = k._bw_state(X.device)
L = _lib.lib()
units = (N // 128) * 4
dbg = torch.zeros(units + N // 128 + 8, 16, device="cuda")
for it in range(2):
    dbg.zero_()
    L.vgan_b200_debug_tile_sums(dbg.data_ptr(), 16)
    out = mmd2_eval(X, None, U, weight=0.0, bw_dev=bw, frozen=(it > 0), mults=k._mults, precision="bf16", masked=True)
    torch.cuda.synchronize()
    k._freeze()
L.vgan_b200_debug_tile_sums(None, 0)
d = dbg[:units, :16].cpu().numpy()
d = dbg[:120, :16].cpu().numpy()      # persistent pair kernel: 15 clusters x 8 CTAs
lead = d[0::2]
peer = d[1::2]
R = (N // 256) * (2 * N // 128 / 4) / 15.0     # rounds per cluster
print(f"persistent pair kernel, cycles per round (4 tiles), {R:.0f} rounds per cluster:")
print("  Gram warp: wait_g_ring %.0f  wait_s_empty %.0f   kernel total %.0f" % (lead[:, 0].mean() / R, lead[:, 1].mean() / R, lead[:, 12].mean() / R))
print("  PV warp:   wait_p %.0f  wait_v_ring %.0f" % (lead[:, 10].mean() / R, lead[:, 11].mean() / R))
print("  SM clock %.0f MHz ; tail (syncs + gradient rows) %.0f cycles per round" % (1e3 * lead[:, 12].mean() / lead[:, 13].mean(), d[:, 14].mean() / R))
print("  epilogue:  wait_s_full %.0f  compute %.0f  wait_own_free %.0f  park+sync %.0f  (the TMEM loads are not in these)" % tuple(d[:, i].mean() / R for i in (4, 5, 6, 7)))
