"""Measured parity of the CUDA path against the fp64 oracle, per case and precision -> gpurun_out/parity_r2.json
(copied to profiles/parity_r2.json).  Same helpers as the -m gpu tests (tests/parity_util.py); nothing is asserted
here, the numbers are recorded.  Run on the B200 box:  python tools/gpu_parity_report.py [--quick]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import parity_util as pu  # noqa: E402
from oracle import mmd_oracle as orc  # noqa: E402
from oracle.make_fixtures import CASES  # noqa: E402


def bench_shape_sample():
    """N = 65536, D = 1024 (BASELINE.json configs[3]): 256 gradient rows and one S_xx row tile against the oracle."""
    from vgan_b200 import RBF
    from vgan_b200.loss import mmd2_eval
    N, D = 65536, 1024
    g = torch.Generator().manual_seed(0)
    X = torch.randn(N, D, generator=g)
    p = torch.softmax(torch.randn(N, D, generator=g), 1)
    U = torch.where(p < 1.0 / D, p, torch.ones_like(p))
    Xd, Ud = X.cuda(), U.cuda()
    for prec in ("bf16", "bf16x3"):
        kernel = RBF()
        bw, frozen = kernel._bw_state(Xd.device)
        r0, r1 = 33024, 33280
        out = mmd2_eval(Xd, None, Ud, weight=0.0, bw_dev=bw, frozen=frozen, mults=kernel._mults, precision=prec,
                        masked=True, row_begin=r0, row_end=r1, xx_part=r0 // 128, xx_nparts=10 ** 6)
        kernel._freeze()
        torch.cuda.synchronize()
        s = out["sums"].cpu().numpy().copy()
        dU = out["dY"].double().cpu()
        bwv = float(kernel.bandwidth)
        t0 = time.time()
        want = orc.bench_sample_fp64(X, U, bwv, r0, r1, r0 // 128)
        pu.record(f"bench_shape/N{N}_D{D}/{prec}", N_x=N, N_y=N, D=D, rows=r1 - r0,
                  sxy_rows=abs(s[1] - want["sxy"]) / want["sxy"], syy_rows=abs(s[2] - want["syy"]) / want["syy"],
                  sxx_tile=abs(s[0] - want["sxx_tile"]) / want["sxx_tile"], dU=pu.rel_l2(dU, want["dU"]),
                  ref_fp32_dU=want.get("ref_fp32_dU"), oracle_seconds=time.time() - t0)
        print(prec, pu._records[f"bench_shape/N{N}_D{D}/{prec}"], flush=True)


def main():
    from vgan_b200 import MMDLossConstrained, RBF
    quick = "--quick" in sys.argv
    for name in CASES:
        for prec in ("bf16", "bf16x3", "simt"):
            err, _ = pu.measure_golden(name, prec, RBF, MMDLossConstrained)
            print(name, prec, {k: f"{v:.2e}" for k, v in err.items()}, flush=True)
    shapes = [(4096, 256), (2048, 1024), (1024, 784), (4096, 64), (640, 1280), (512, 2048), (1000, 512), (4096, 32),
              (2048, 128)]
    for N, D in shapes:
        err, _ = pu.measure_masked(N, D, "bf16", RBF, MMDLossConstrained, seed=N + D)
        print(N, D, {k: f"{v:.2e}" for k, v in err.items()}, flush=True)
    pu.flush()
    if not quick:
        bench_shape_sample()
    pu.flush()
    print("wrote", pu.REPORT)


if __name__ == "__main__":
    main()
