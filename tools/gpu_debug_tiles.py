import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vgan_b200 import RBF, _lib
from vgan_b200.loss import mmd2_eval
from oracle import mmd_oracle as orc

N, D = 4096, 64
rs = np.random.RandomState(0)
X = rs.randn(N, D).astype(np.float32); U = rs.rand(N, D).astype(np.float32)
Z = np.vstack([X, U * X]).astype(np.float64)
bw = orc.bandwidth_direct(Z)
T = N // 128
Xd, Ud = torch.tensor(X, device="cuda"), torch.tensor(U, device="cuda")
k = RBF(); bwd, _ = k._bw_state(Xd.device)
L = _lib.lib()
units = T + T
dbg = torch.zeros(units, 2 * T, device="cuda")
L.vgan_b200_debug_tile_sums(dbg.data_ptr(), 2 * T)
out = mmd2_eval(Xd, None, Ud, weight=0.0, bw_dev=bwd, frozen=False, mults=k._mults, precision="bf16",
                need_dX=False, need_dY=False, masked=True)
torch.cuda.synchronize()
L.vgan_b200_debug_tile_sums(None, 0)
got = dbg.cpu().numpy()
# reference per-tile sums (fp64) of the rounded operand
Zt = torch.tensor(Z)
Kfull = orc.rbf_matrix(Zt, bw).numpy()
ref = Kfull.reshape(2 * T, 128, 2 * T, 128).sum((1, 3))     # [row tile, col tile]
# unit order: T Y-row units (row tile T+u, all 2T col tiles), then T sym units (row tile x, cols x..T-1)
bad = []
for u_ in range(T):
    for n in range(2 * T):
        r = ref[T + u_, n]; g = got[u_, n]
        if abs(g - r) > 2e-3 * r: bad.append(("Y", u_, n, g, r))
for x in range(T):
    for n in range(T - x):
        r = ref[x, x + n]; g = got[T + x, n]
        if abs(g - r) > 2e-3 * r: bad.append(("S", x, n, g, r))
print("bad tiles:", len(bad))
for b in bad[:60]: print(b)
