"""Per-role wait timers of the gradient (pair-group) kernel: cycles spent waiting on each barrier, per round of CS column
tiles.  Builds and loads the debug library with TIMERS=1 (the product build compiles the clock reads out)."""
import ctypes as C, os, subprocess, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if "--no-build" not in sys.argv:      # (--no-build: libvgan_b200_dbg.so was built with DEBUG=1 TIMERS=1 before the snapshot)
    subprocess.run(["make", "-C", os.path.join(ROOT, "vgan_b200", "csrc"), "-j8", "DEBUG=1", "TIMERS=1"], check=True,
                   capture_output=True, env=dict(os.environ, PATH="/usr/local/cuda/bin:" + os.environ["PATH"]))
sys.argv = [a for a in sys.argv if a != "--no-build"]
from vgan_b200 import _lib
L = C.CDLL(_lib.DBG_LIB_PATH); _lib._declare(L); _lib._lib = L
L.vgan_b200_debug_tile_sums.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
from vgan_b200 import RBF
from vgan_b200.loss import mmd2_eval
N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
D = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
torch.manual_seed(0)
X = torch.randn(N, D, device="cuda"); U = torch.rand(N, D, device="cuda")
k = RBF(); bw, _ = k._bw_state(X.device)
stride = 48
dbg = torch.zeros(256, stride, device="cuda")
for it in range(2):
    dbg.zero_()
    L.vgan_b200_debug_tile_sums(_lib.ctx(0), dbg.data_ptr(), stride)
    out = mmd2_eval(X, None, U, weight=0.0, bw_dev=bw, frozen=(it > 0), mults=k._mults, precision="bf16", masked=True)
    torch.cuda.synchronize()
    k._freeze()
L.vgan_b200_debug_tile_sums(_lib.ctx(0), None, 0)
d = dbg.cpu().numpy()
used = d[:, 12] > 0
lead = d[used]
allc = d[(d[:, 4:8].sum(1) > 0)]
groups = len(lead) // 4
R = (N // 256) * (2 * N // 128 / 4) / groups     # rounds per group
print(f"{groups} groups, cycles per round (4 tiles), {R:.0f} rounds per group:")
print("  (raw, cycles per group) kernel total %.0f  tail %.0f  gram waits %.0f + %.0f  epilogue wait_s_full %.0f compute %.0f" % (
    lead[:, 12].mean(), allc[:, 14].mean(), lead[:, 0].mean(), lead[:, 1].mean(), allc[:, 4].mean(), allc[:, 5].mean()))
print("  tail parts (cycles per group): barrier #1 %.0f  row sums + barrier #2 %.0f  gradient rows %.0f" % tuple(allc[:, i].mean() for i in (16, 17, 18)))
print("  gradient rows (cycles per group): TMEM -> shared %.0f  rows %.0f  column sums %.0f" % tuple(allc[:, i].mean() for i in (20, 21, 22)))
print("  Gram warp: wait_g_ring %.0f  wait_s_empty %.0f   kernel total %.0f" % (lead[:, 0].mean() / R, lead[:, 1].mean() / R, lead[:, 12].mean() / R))
print("  PV warp:   wait_p %.0f  wait_v_ring %.0f" % (lead[:, 10].mean() / R, lead[:, 11].mean() / R))
print("  SM clock %.0f MHz ; tail (syncs + gradient rows) %.0f cycles per round" % (1e3 * lead[:, 12].mean() / lead[:, 13].mean(), allc[:, 14].mean() / R))
print("  epilogue:  wait_s_full %.0f  compute %.0f  wait_own_free %.0f  park+sync %.0f  (the TMEM loads are not in these)" % tuple(allc[:, i].mean() / R for i in (4, 5, 6, 7)))
