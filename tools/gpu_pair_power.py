"""Is the gradient kernel bound by the SMs it occupies or by the power cap?  Times K1 (pair kernel + S_xx kernel) at the
bench shape with the persistent grid limited to fewer clusters (vgan_b200_set_cluster_limit) and samples the SM clock:
SM-bound => time ~ 1/clusters at a constant clock; power-bound => fewer clusters run at a higher clock."""
import json, os, subprocess, sys, threading, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import make_inputs, ClockSampler
from vgan_b200 import MMDLossConstrained, RBF, _lib
from vgan_b200.dist import sharded_masked_step

def main():
    N, D = 65536, 1024
    X, U = make_inputs(N, D)
    X, U = X.cuda(), U.cuda()
    f = MMDLossConstrained(10.0, kernel=RBF(), precision="bf16")
    L, ctx = _lib.lib(), _lib.ctx(0)
    for _ in range(3):
        sharded_masked_step(f, X, U)
    torch.cuda.synchronize()
    for limit in [int(a) for a in sys.argv[1:]] or [0, 14, 12, 10, 8, 0]:
        _lib.check(L.vgan_b200_set_cluster_limit(ctx, limit), "limit")
        sharded_masked_step(f, X, U)
        torch.cuda.synchronize()
        s = ClockSampler(0); s.start()
        evs = []
        t0 = time.perf_counter()
        while time.perf_counter() - t0 < 1.5:
            evs.append(sharded_masked_step(f, X, U)["k1_events"])
            torch.cuda.synchronize()
        s.stop_flag = True
        ms = float(np.median([a.elapsed_time(b) for a, b in evs]))
        print(json.dumps({"cluster_limit": limit, "k1_ms": round(ms, 3), "steps": len(evs), "clocks": s.result(),
                          "ms_x_clusters": round(ms * (limit or 15), 1)}), flush=True)
    L.vgan_b200_set_cluster_limit(ctx, 0)

if __name__ == "__main__":
    main()
