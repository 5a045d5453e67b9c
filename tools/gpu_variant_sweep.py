"""Pipeline-shape sweep of the CTA-pair K1 kernel: time per evaluation and agreement with variant 0."""
import os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vgan_b200 import RBF, _lib
from vgan_b200.loss import mmd2_eval
N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
D = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
variants = [int(v) for v in sys.argv[3].split(",")] if len(sys.argv) > 3 else list(range(7))
torch.manual_seed(0)
X = torch.randn(N, D, device="cuda"); U = torch.rand(N, D, device="cuda")
k = RBF(); bw, _ = k._bw_state(X.device)
L = _lib.lib()
ref = None
for v in variants:
    L.vgan_b200_debug_variant(v)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ts = []
    for it in range(5):
        out = mmd2_eval(X, None, U, weight=0.0, bw_dev=bw, frozen=(k.bandwidth is not None), mults=k._mults,
                        precision="bf16", masked=True, k1_events=ev)
        torch.cuda.synchronize(); k._freeze()
        if it >= 2: ts.append(ev[0].elapsed_time(ev[1]))
    sums = out["sums"].clone(); dY = out["dY"].clone()
    if ref is None: ref = (sums, dY)
    err_s = float(((sums[:3] - ref[0][:3]).abs() / ref[0][:3].abs()).max())
    err_g = float((dY - ref[1]).norm() / ref[1].norm())
    print(json.dumps({"variant": v, "N": N, "D": D, "k1_ms": round(sum(ts) / len(ts), 3),
                      "sum_rel_vs_v0": err_s, "grad_rel_vs_v0": err_g}), flush=True)
L.vgan_b200_debug_variant(0)
