"""Gradient kernel alone at the bench shape with the persistent grid capped to 6..15 clusters (48..120 SMs): does the
throughput scale with the SM count, or does the power cap eat it?  Prints ms, SM-ms and the in-run SM clock."""
import os, sys, subprocess, threading, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vgan_b200 import RBF, _lib
from vgan_b200.loss import mmd2_eval
N, D = 65536, 1024
torch.manual_seed(0)
X = torch.randn(N, D, device="cuda"); U = torch.rand(N, D, device="cuda")
k = RBF(); bw, _ = k._bw_state(X.device)
L = _lib.lib()

def clock():
    out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits", "-i", "0"],
                         capture_output=True, text=True).stdout.strip().split(",")
    return float(out[0]), float(out[1])

for cap in (15, 12, 9, 6, 15):
    L.vgan_b200_debug_variant(cap)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ts, clk = [], []
    for it in range(10):
        mmd2_eval(X, None, U, weight=0.0, bw_dev=bw, frozen=(k.bandwidth is not None), mults=k._mults,
                  precision="bf16", masked=True, k1_events=ev, row_begin=0, row_end=N, xx_part=10**6 - 1, xx_nparts=10**6)
        if it >= 3: clk.append(clock())
        torch.cuda.synchronize(); k._freeze()
        if it >= 3: ts.append(ev[0].elapsed_time(ev[1]))
    t = sum(ts) / len(ts)
    print(f"clusters={cap:2d} ({cap*8:3d} SMs): K1 {t:7.3f} ms   SM-ms {t*cap*8:8.1f}   clock {sum(c[0] for c in clk)/len(clk):6.0f} MHz  power {sum(c[1] for c in clk)/len(clk):5.0f} W", flush=True)
L.vgan_b200_debug_variant(0)
