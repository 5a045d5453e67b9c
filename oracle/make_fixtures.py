"""Generate tests/golden/*.npz by running the REFERENCE's own loss module in fp64.

TEST INFRASTRUCTURE ONLY.  Run in the authoring container (needs /root/reference):

    python oracle/make_fixtures.py

Imports /root/reference/src/models/Mmd_loss_constrained.py (it depends on torch only)
and, per case, builds seeded inputs, a FRESH ``RBF()`` (the reference's default-argument
RBF is shared process-wide and freezes its bandwidth, Mmd_loss_constrained.py:16-22,35),
runs forward + autograd backward in float64 and stores inputs' recipe + outputs.
Inputs are re-created from the seed by ``build_case`` (also used by the tests), so the
.npz files hold only outputs (and a checksum of the inputs).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(os.path.dirname(HERE), "tests", "golden")

# name -> (Nx, Ny, D, weight, kind)
CASES = {
    "tiny_8x4":        (8, 8, 4, 10.0, "mask"),
    "ragged_37x5":     (37, 37, 5, 10.0, "mask"),
    "tile_128x16":     (128, 128, 16, 0.0, "mask"),
    "cfg1_500x20":     (500, 500, 20, 10.0, "generator"),
    "kl_500x1":        (500, 500, 1, 0.0, "free"),
    "kl_256x49":       (256, 256, 49, 0.0, "free"),
    "unequal_200_328": (200, 328, 24, 0.0, "free"),
    "dups_96x12":      (96, 96, 12, 10.0, "dups"),
    "cov_300x10":      (300, 300, 10, 10.0, "cov"),
    "wide_256x784":    (256, 256, 784, 10.0, "mnist"),
    "wide_384x1024":   (384, 384, 1024, 0.0, "mask"),
}


def upper_softmax(h: np.ndarray) -> np.ndarray:
    """Generator.py:18-22 — softmax over features, entries >= 1/D replaced by 1."""
    h = h - h.max(1, keepdims=True)
    p = np.exp(h)
    p /= p.sum(1, keepdims=True)
    D = h.shape[1]
    return np.where(p < 1.0 / D, p, 1.0)


def build_case(name: str):
    """Seeded inputs (float64 numpy): X [Nx,D], Y [Ny,D], U [Nx,D] or None, weight, kind.

    kind 'mask'/'generator'/'mnist'/'dups'/'cov': Y = U * X (the VGAN_no_kl call,
    vgan.py:614-616).  kind 'free': X and Y independent (the VGAN encodings call,
    vgan.py:283), U only feeds the zero-weight penalty."""
    Nx, Ny, D, weight, kind = CASES[name]
    seed = sum(ord(ch) for ch in name) * 7919 % (2 ** 31)
    rs = np.random.RandomState(seed)
    if kind == "cov":
        A = rs.randn(D, D)
        X = rs.randn(Nx, D) @ A.T * 0.5 + rs.randn(D)
    elif kind == "mnist":
        X = rs.rand(Nx, D) * (rs.rand(Nx, D) < 0.2)
    else:
        X = rs.randn(Nx, D)
    if kind == "free":
        Y = rs.randn(Ny, D) * 1.3 + 0.25
        U = rs.rand(Nx, max(D, 4))
        return X, Y, U, weight, kind
    if kind == "generator" or kind == "mnist":
        U = upper_softmax(rs.randn(Nx, D) * 1.5)
    else:
        U = rs.rand(Nx, D)
    if kind == "dups":
        X[1] = X[0]
        X[7] = X[5]
        U[1] = U[0]
        U[:, 3] = 1.0      # tied column maxima -> lowest row index takes the gradient
        U[:, 4] = 0.5
    Y = U * X
    return X, Y, U, weight, kind


def load_reference():
    sys.path.insert(0, "/root/reference")
    from src.models.Mmd_loss_constrained import MMDLossConstrained, RBF  # type: ignore
    return MMDLossConstrained, RBF


def run_reference(name: str):
    MMD, RBF = load_reference()
    X, Y, U, weight, kind = build_case(name)
    Xt = torch.tensor(X, dtype=torch.float64, requires_grad=True)
    Ut = torch.tensor(U, dtype=torch.float64, requires_grad=True)
    if kind == "free":
        Yt = torch.tensor(Y, dtype=torch.float64, requires_grad=True)
    else:
        Yt = Ut * Xt.detach()
    loss_fn = MMD(weight, kernel=RBF())
    loss_fn.device = torch.device("cpu")
    loss_fn.kernel.bandwidth_multipliers = loss_fn.kernel.bandwidth_multipliers.to("cpu")
    loss = loss_fn(Xt, Yt, Ut)
    if kind == "free":
        Yt.retain_grad()
    loss.backward()
    out = dict(loss=float(loss), bandwidth=float(loss_fn.bandwidth),
               mults=loss_fn.bandwidth_multipliers.double().numpy(),
               dX=Xt.grad.numpy(), dU=(Ut.grad.numpy() if Ut.grad is not None else np.zeros_like(U)),
               in_checksum=np.array([X.sum(), Y.sum(), U.sum()]))
    if kind == "free":
        out["dY"] = Yt.grad.numpy()
    # block sums from the reference's own kernel matrix
    with torch.no_grad():
        K = loss_fn.kernel(torch.vstack([Xt, Yt]))
        nx = X.shape[0]
        out["sxx"] = float(K[:nx, :nx].sum())
        out["sxy"] = float(K[:nx, nx:].sum())
        out["syy"] = float(K[nx:, nx:].sum())
    # second call on the SAME module with shifted data: bandwidth must stay frozen
    with torch.no_grad():
        loss2 = loss_fn(Xt * 1.5, Yt * 1.5, Ut)
        out["loss_frozen_bw_scaled"] = float(loss2)
        out["bandwidth_after_second_call"] = float(loss_fn.bandwidth)
    return out


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    for name in CASES:
        out = run_reference(name)
        Nx, Ny, D, weight, kind = CASES[name]
        if Nx * D > 64 * 1024:      # keep fixtures small: first 24 rows + column sums
            for key in ("dX", "dU", "dY"):
                if key in out:
                    g = out.pop(key)
                    out[key + "_head"] = g[:24].copy()
                    out[key + "_colsum"] = g.sum(0)
                    out[key + "_l2"] = np.array(np.sqrt((g * g).sum()))
        np.savez_compressed(os.path.join(GOLDEN, name + ".npz"), **out)
        print(f"{name:18s} loss={out['loss']:.12g} bw={out['bandwidth']:.9g}")


if __name__ == "__main__":
    main()
