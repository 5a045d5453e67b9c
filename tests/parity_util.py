"""Shared measurement helpers of the GPU parity tests and tools/gpu_parity_report.py: run one case through the
public modules (ctypes -> C ABI -> CUDA kernels), compare with the fp64 oracle and return the MEASURED errors, so
that the tests assert on them and the report (profiles/parity_r2.json) records them.  TEST INFRASTRUCTURE.

Error measures: relative error of the three kernel sums, relative L2 error of each gradient
array.  ``yardstick`` = the same errors of the reference formula evaluated by torch in fp32 on the same device
(oracle/torch_eager.py, pinned to the golden vectors) — what the reference itself delivers at its own precision."""
import json
import os

import numpy as np
import torch

from oracle import mmd_oracle as orc
from oracle import torch_eager
from oracle.make_fixtures import build_case

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REPORT = os.path.join(ROOT, "gpurun_out", "parity_r2.json")

_records = {}


def record(key: str, **vals):
    _records.setdefault(key, {}).update({k: (float(v) if v is not None else None) for k, v in vals.items()})


def flush():
    if not _records:
        return
    os.makedirs(os.path.dirname(REPORT), exist_ok=True)
    old = {}
    if os.path.exists(REPORT):
        try:
            with open(REPORT) as f:
                old = json.load(f)
        except Exception:
            old = {}
    old.update(_records)
    with open(REPORT, "w") as f:
        json.dump(old, f, indent=1, sort_keys=True)


def rel_l2(got, want) -> float:
    got = torch.as_tensor(got).double().cpu()
    want = torch.as_tensor(want).double().cpu()
    return float((got - want).norm() / want.norm())


def eager_fp32_yardstick(X, Y, U, weight, kind, ref):
    """Reference formula in fp32 torch on the GPU (same inputs): its own error against the fp64 oracle."""
    n = X.shape[0] + Y.shape[0]
    if n * n * 4 * 8 > 8e9:
        return {}
    Xd = torch.tensor(X, dtype=torch.float32, device="cuda", requires_grad=True)
    Ud = torch.tensor(U, dtype=torch.float32, device="cuda", requires_grad=True)
    if kind == "free":
        Yd = torch.tensor(Y, dtype=torch.float32, device="cuda", requires_grad=True)
    else:
        Yd = Ud * Xd.detach()
    f = torch_eager.MMDLossConstrained(weight, kernel=torch_eager.RBF(device="cuda"))
    loss = f(Xd, Yd, Ud)
    loss.backward()
    out = {"ref_fp32_dX": rel_l2(Xd.grad, ref["dX"])}
    if kind == "free":
        out["ref_fp32_dY"] = rel_l2(Yd.grad, ref["dY"])
    else:
        out["ref_fp32_dU"] = rel_l2(Ud.grad, ref["dY"] * torch.tensor(X) + ref["dU_penalty"])
    return out


def measure_case(X, Y, U, weight, kind, prec, RBF_cls, MMD_cls, yardstick=True):
    """One forward+backward through the nn.Module surface; returns (errors, objects) with errors =
    dict(sums, loss_scaled, dX, dY | dU, bandwidth)."""
    Xd = torch.tensor(X, dtype=torch.float32, device="cuda", requires_grad=True)
    Ud = torch.tensor(U, dtype=torch.float32, device="cuda", requires_grad=True)
    loss_fn = MMD_cls(weight, kernel=RBF_cls(), precision=prec)
    if kind == "free":
        Yd = torch.tensor(Y, dtype=torch.float32, device="cuda", requires_grad=True)
    else:
        Yd = Ud * Xd.detach()
    loss = loss_fn(Xd, Yd, Ud)
    loss.backward()
    torch.cuda.synchronize()
    ref = orc.loss_and_grads(X, Y, U, weight, need_dX=True)
    s = loss_fn.last_sums.cpu().numpy()
    nx, ny = X.shape[0], Y.shape[0]
    scale = ref["sxx"] / nx ** 2 + 2 * ref["sxy"] / (nx * ny) + ref["syy"] / ny ** 2
    err = {
        "sums": max(abs(float(g) - ref[k]) / ref[k] for g, k in zip(s[:3], ("sxx", "sxy", "syy"))),
        "loss_scaled": abs(float(loss.detach()) - ref["loss"]) / scale,
        "bandwidth": abs(float(loss_fn.bandwidth) - ref["bandwidth"]) / ref["bandwidth"],
        "dX": rel_l2(Xd.grad, ref["dX"]),
    }
    if kind == "free":
        err["dY"] = rel_l2(Yd.grad, ref["dY"])
    else:
        err["dU"] = rel_l2(Ud.grad, ref["dY"] * torch.tensor(X) + ref["dU_penalty"])
    if yardstick:
        err.update(eager_fp32_yardstick(X, Y, U, weight, kind, ref))
    objs = dict(loss=loss, loss_fn=loss_fn, Xd=Xd, Yd=Yd, Ud=Ud, ref=ref, sums=s)
    return err, objs


def measure_golden(name, prec, RBF_cls, MMD_cls, yardstick=True):
    X, Y, U, weight, kind = build_case(name)
    err, objs = measure_case(X, Y, U, weight, kind, prec, RBF_cls, MMD_cls, yardstick)
    record(f"golden/{name}/{prec}", N_x=X.shape[0], N_y=Y.shape[0], D=X.shape[1], **err)
    return err, objs


def measure_masked(N, D, prec, RBF_cls, MMD_cls, seed, weight=0.0, tag="medium", yardstick=True):
    rs = np.random.RandomState(seed)
    X = rs.randn(N, D).astype(np.float32)
    U = rs.rand(N, D).astype(np.float32)
    err, objs = measure_case(X.astype(np.float64), (U * X).astype(np.float64), U.astype(np.float64), weight, "mask",
                             prec, RBF_cls, MMD_cls, yardstick)
    record(f"{tag}/N{N}_D{D}/{prec}", N_x=N, N_y=N, D=D, **err)
    return err, objs
