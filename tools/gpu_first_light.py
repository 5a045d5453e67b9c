"""Diagnostics run on the GPU box: probe the UMMA descriptors, then every golden case in both
precisions, printing errors instead of asserting.  (Uses the oracle as checker only.)"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vgan_b200 import _lib, MMDLossConstrained, RBF  # noqa: E402
from oracle import mmd_oracle as orc  # noqa: E402
from oracle.make_fixtures import CASES, GOLDEN, build_case  # noqa: E402


def probe():
    L = _lib.lib()
    torch.manual_seed(0)
    A = torch.randn(128, 128, device="cuda").bfloat16().contiguous()
    B = torch.randn(128, 128, device="cuda").bfloat16().contiguous()
    S = torch.zeros(128, 128, device="cuda")
    O = torch.zeros(128, 128, device="cuda")
    _lib.check(L.vgan_b200_umma_probe(A.data_ptr(), B.data_ptr(), S.data_ptr(), O.data_ptr(),
                                      torch.cuda.current_stream().cuda_stream), "probe")
    torch.cuda.synchronize()
    S_ref = A.float() @ B.float().T
    O_ref = S.bfloat16().float() @ B.float()
    es = float((S - S_ref).abs().max() / S_ref.abs().max())
    eo = float((O - O_ref).abs().max() / O_ref.abs().max())
    print(f"probe: S rel-max-err {es:.3e}   O rel-max-err {eo:.3e}")
    if eo > 1e-2:
        # which alternative would match? helps to localise a descriptor mistake
        alt = {"O=P@B^T": S.bfloat16().float() @ B.float().T, "O=P^T@B": S.bfloat16().float().T @ B.float()}
        for k, v in alt.items():
            print("   ", k, float((O - v).abs().max() / v.abs().max()))
    return es < 1e-2 and eo < 1e-2


def case(name, prec):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    X, Y, U, weight, kind = build_case(name)
    dev = "cuda"
    Xd = torch.tensor(X, dtype=torch.float32, device=dev, requires_grad=True)
    Ud = torch.tensor(U, dtype=torch.float32, device=dev, requires_grad=True)
    loss_fn = MMDLossConstrained(weight, kernel=RBF(), precision=prec)
    if kind == "free":
        Yd = torch.tensor(Y, dtype=torch.float32, device=dev, requires_grad=True)
    else:
        Yd = Ud * Xd.detach()
    t0 = time.time()
    loss = loss_fn(Xd, Yd, Ud)
    loss.backward()
    torch.cuda.synchronize()
    dt = time.time() - t0
    s = loss_fn.last_sums.cpu().numpy()
    es = max(abs(s[0] - g["sxx"]) / g["sxx"], abs(s[1] - g["sxy"]) / g["sxy"], abs(s[2] - g["syy"]) / g["syy"])
    ref = orc.loss_and_grads(X, Y, U, weight, need_dX=True)
    dX = Xd.grad.double().cpu()
    ex = float((dX - ref["dX"]).norm() / ref["dX"].norm())
    if kind == "free":
        gy = Yd.grad.double().cpu()
        ey = float((gy - ref["dY"]).norm() / ref["dY"].norm())
    else:
        dU_ref = ref["dY"] * torch.tensor(X) + ref["dU_penalty"]
        gy = Ud.grad.double().cpu()
        ey = float((gy - dU_ref).norm() / dU_ref.norm())
    ebw = abs(float(loss_fn.bandwidth) - float(g["bandwidth"])) / float(g["bandwidth"])
    print(f"{name:18s} {prec}: loss {float(loss):.6f} ref {float(g['loss']):.6f} | sums rel {es:.2e} | "
          f"bw rel {ebw:.1e} | dX relL2 {ex:.2e} | dY/dU relL2 {ey:.2e} | {dt*1e3:.1f} ms")


if __name__ == "__main__":
    print(torch.cuda.get_device_name(0))
    ok = probe()
    for prec in ("fp32", "bf16"):
        if prec == "bf16" and not ok:
            print("probe failed: still running bf16 cases for information")
        for name in CASES:
            try:
                case(name, prec)
            except Exception as e:  # noqa: BLE001
                print(f"{name} {prec}: EXCEPTION {type(e).__name__}: {e}")
                if "CUDA" in str(e) or "cuda" in str(e):
                    raise
