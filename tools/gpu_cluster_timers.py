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
if not os.environ.get("VGAN_B200_NO_PAIR"):
    d = d[256:]          # the S_xx side kernel writes its tile sums into the first rows of the same buffer
    lead = d[d[:, 0] > 0]
    R = 2 * N // 128 / 4
    print("pair leaders, cycles per round (4 tiles):")
    print("  Gram warp: total %.0f  wait_g_ring %.0f  wait_s_empty %.0f" % tuple(lead[:, i].mean() / R for i in (0, 2, 3)))
    print("  PV warp:   total %.0f  wait_p %.0f  wait_v_ring %.0f" % tuple(lead[:, i].mean() / R for i in (8, 1, 9)))
names = ["mma_total", "mma_wait_pfull", "mma_wait_ring", "mma_wait_sempty", "epi_wait_sfull", "epi_compute", "epi_wait_slot", "epi_publish"]
tiles = 2 * N // 128
print(f"N={N} tiles={tiles} rounds={tiles//4}; cycles per ROUND (4 tiles), mean over CTAs:")
for i, n in enumerate(names):
    print(f"  {n:18s} {d[:, i].mean() / (tiles / 4):10.0f}")
