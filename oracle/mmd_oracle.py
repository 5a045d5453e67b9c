"""CPU oracle for V-GAN's squared-MMD hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``vgan_b200/`` may import this file; only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs do, and there only as the checker / the timed CPU baseline.

It is a fp64 restatement (numpy / torch-CPU, blockwise over column tiles so that it
runs at any N without the reference's ``[5, 2N, 2N]`` intermediate) of

* ``RBF.get_bandwidth``      /root/reference/src/models/Mmd_loss_constrained.py:16-22
* ``RBF.forward``            /root/reference/src/models/Mmd_loss_constrained.py:24-26
* ``MMDLossConstrained.forward`` + its autograd backward
                             /root/reference/src/models/Mmd_loss_constrained.py:42-50
* the masking ``Y = U * X``  /root/reference/src/vgan.py:616
* ``torch_two_sample.MMDStatistic`` (third party, josipd/torch-two-sample, UNPINNED
  in the reference: README.md:45-51 clones default-branch HEAD; absent from this
  machine).  Restated from its published algorithm; call sites
  /root/reference/src/vgan.py:419-428, 696-705.  **parity unpinned** for this one
  function group (``mmd_statistic`` / ``permutation_pval``): the reference holds no
  golden vectors for it and the package cannot be imported here.

Parity pin for everything else: ``tests/golden/*.npz`` were produced by importing the
reference's own ``Mmd_loss_constrained.py`` in fp64 (``oracle/make_fixtures.py``) and
``tests/test_oracle.py`` checks this restatement against them (and, when
``/root/reference`` is present, against the live module).
"""
from __future__ import annotations

import math
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

DEFAULT_N_KERNELS = 5
DEFAULT_MUL_FACTOR = 2.0


def bandwidth_multipliers(n_kernels: int = DEFAULT_N_KERNELS,
                          mul_factor: float = DEFAULT_MUL_FACTOR) -> np.ndarray:
    """``mul_factor ** (arange(n) - n // 2)``  (Mmd_loss_constrained.py:12-13)."""
    return mul_factor ** (np.arange(n_kernels) - n_kernels // 2).astype(np.float64)


def _f64(a) -> torch.Tensor:
    if isinstance(a, torch.Tensor):
        return a.detach().to("cpu", torch.float64)
    return torch.as_tensor(np.asarray(a), dtype=torch.float64)


def scales(bandwidth: float, mults) -> list:
    """s_m = bandwidth * mult_m as the reference forms it: a 0-dim (fp64 or fp32) bandwidth
    times the float32 multiplier VECTOR promotes to float32 (Mmd_loss_constrained.py:26),
    so every s_m is rounded to fp32 even when the data is fp64."""
    return [float(np.float32(bandwidth) * np.float32(m)) for m in mults]


def _sqdist(A: torch.Tensor, B: torch.Tensor) -> torch.Tensor:
    """Squared euclidean distances, clamped at 0 (what ``cdist(...)**2`` yields)."""
    na = (A * A).sum(1)[:, None]
    nb = (B * B).sum(1)[None, :]
    return (na + nb - 2.0 * (A @ B.T)).clamp_min_(0.0)


def bandwidth_direct(Z) -> float:
    """sum_ij d2_ij / (n^2 - n) over the stacked sample (Mmd_loss_constrained.py:16-22)."""
    Z = _f64(Z)
    n = Z.shape[0]
    total = 0.0
    for s in range(0, n, 2048):
        total += float(_sqdist(Z[s:s + 2048], Z).sum())
    return total / (n * n - n)


def bandwidth_closed_form(Z) -> float:
    """Same value in O(nD): sum_ij |zi-zj|^2 = 2n sum|z|^2 - 2 |sum z|^2."""
    Z = _f64(Z)
    n = Z.shape[0]
    sumsq = float((Z * Z).sum())
    s = Z.sum(0)
    return (2.0 * n * sumsq - 2.0 * float(s @ s)) / (n * n - n)


def rbf_matrix(Z, bandwidth: float, mults: Optional[Sequence[float]] = None) -> torch.Tensor:
    """``RBF.forward`` with a given (frozen) bandwidth -> [n, n] (small n only)."""
    Z = _f64(Z)
    mults = bandwidth_multipliers() if mults is None else np.asarray(mults, np.float64)
    d2 = _sqdist(Z, Z)
    K = torch.zeros_like(d2)
    for sm in scales(bandwidth, mults):
        K += torch.exp(-d2 / sm)
    return K


def colmax_penalty(U, weight: float) -> Tuple[float, torch.Tensor]:
    """``weight * mean_d(1 - max_i U[i,d])`` and its gradient.

    Mmd_loss_constrained.py:50 uses ``torch.topk(U, 1, 0)``; on ties CPU torch routes the
    whole gradient to the lowest row index (probed), pinned here.
    """
    U = _f64(U)
    D = U.shape[1]
    vals, idx = U.max(dim=0)
    # torch.max returns the first maximal index on CPU; make it explicit.
    idx = torch.argmax((U == vals[None, :]).to(torch.int8), dim=0)
    g = torch.zeros_like(U)
    g[idx, torch.arange(D)] = -float(weight) / D
    return float(weight) * float((1.0 - vals).mean()), g


def mmd2_terms(X, Y, bandwidth: float, mults: Optional[Sequence[float]] = None,
               need_dX: bool = True, block: int = 2048, eps: float = 0.0):
    """Blockwise S_xx, S_xy, S_yy and dL/dX, dL/dY of the biased MMD^2.

    L = S_xx/Nx^2 - 2 S_xy/(Nx Ny) + S_yy/Ny^2      (Mmd_loss_constrained.py:46-50)
    With signed weights c = [+1/Nx]*Nx ++ [-1/Ny]*Ny over the stacked Z = [X; Y]
    (Mmd_loss_constrained.py:43):  L = sum_ij c_i c_j k_ij  and
      dL/dz_i = 4 c_i ( sum_j c_j w_ij z_j  -  (sum_j c_j w_ij) z_i ),
      k = sum_m exp(-d2/s_m),  w = sum_m exp(-d2/s_m)/s_m,  s_m = bandwidth * mult_m.
    ``eps`` is added to d2 (torch_two_sample's pdist uses 1e-5; 0 for the loss).
    """
    X = _f64(X)
    Y = _f64(Y)
    mults = bandwidth_multipliers() if mults is None else np.asarray(mults, np.float64)
    s = scales(bandwidth, mults)
    Nx, Ny = X.shape[0], Y.shape[0]
    Z = torch.cat([X, Y], 0)
    c = torch.cat([torch.full((Nx,), 1.0 / Nx, dtype=torch.float64),
                   torch.full((Ny,), -1.0 / Ny, dtype=torch.float64)])
    n = Nx + Ny
    sums = {"xx": 0.0, "xy": 0.0, "yy": 0.0}
    G = torch.zeros_like(Z)
    row0 = 0 if need_dX else Nx
    # rows [row0, n) need gradients; S_xx additionally needs the X x X block.
    for i0 in range(0, n, block):
        i1 = min(i0 + block, n)
        Zi = Z[i0:i1]
        want_grad = i1 > row0
        O = torch.zeros_like(Zi)
        r = torch.zeros(i1 - i0, dtype=torch.float64)
        for j0 in range(0, n, block):
            j1 = min(j0 + block, n)
            if not want_grad and j0 >= Nx:
                continue
            d2 = _sqdist(Zi, Z[j0:j1]) + eps
            k = torch.zeros_like(d2)
            w = torch.zeros_like(d2)
            for sm in s:
                e = torch.exp(-d2 / sm)
                k += e
                w += e / sm
            # block sums, split at the X/Y boundary on both axes
            ri = torch.arange(i0, i1) < Nx
            cj = torch.arange(j0, j1) < Nx
            kx = k[ri]
            ky = k[~ri]
            sums["xx"] += float(kx[:, cj].sum())
            sums["xy"] += float(ky[:, cj].sum())      # = K_xy by symmetry; Y rows see every column
            sums["yy"] += float(ky[:, ~cj].sum())
            if want_grad:
                P = w * c[None, j0:j1]
                O += P @ Z[j0:j1]
                r += P.sum(1)
        if want_grad:
            G[i0:i1] = 4.0 * c[i0:i1, None] * (O - r[:, None] * Zi)
    dX = G[:Nx] if need_dX else None
    dY = G[Nx:]
    return sums["xx"], sums["xy"], sums["yy"], dX, dY


def mmd2_from_sums(sxx: float, sxy: float, syy: float, Nx: int, Ny: int) -> float:
    return sxx / (Nx * Nx) - 2.0 * sxy / (Nx * Ny) + syy / (Ny * Ny)


def loss_and_grads(X, Y, U, weight: float, bandwidth: Optional[float] = None,
                   mults: Optional[Sequence[float]] = None, need_dX: bool = True,
                   block: int = 2048):
    """Full ``MMDLossConstrained.forward(X, Y, U)`` value and gradients w.r.t. X, Y, U.

    ``bandwidth=None`` reproduces the first call of a fresh ``RBF()`` (computes it from
    the stacked input and would freeze it).  The U gradient here is only the penalty's
    (the ``Y = U*X`` chain is applied by the caller: dU += dY * X, vgan.py:616).
    Returns dict(loss, mmd2, bandwidth, sxx, sxy, syy, dX, dY, dU_penalty, penalty).
    """
    Xd, Yd = _f64(X), _f64(Y)
    if bandwidth is None:
        bandwidth = bandwidth_direct(torch.cat([Xd, Yd], 0))
    sxx, sxy, syy, dX, dY = mmd2_terms(Xd, Yd, bandwidth, mults, need_dX, block)
    mmd2 = mmd2_from_sums(sxx, sxy, syy, Xd.shape[0], Yd.shape[0])
    pen, dUp = colmax_penalty(U, weight)
    return dict(loss=mmd2 + pen, mmd2=mmd2, bandwidth=float(bandwidth), sxx=sxx, sxy=sxy,
                syy=syy, dX=dX, dY=dY, dU_penalty=dUp, penalty=pen)


def masked_loss_and_grad(X, U, weight: float, bandwidth: Optional[float] = None,
                         mults: Optional[Sequence[float]] = None, block: int = 2048):
    """The ``VGAN_no_kl`` step: ``loss(X, U*X, U)`` and dL/dU (vgan.py:614-618)."""
    Xd, Ud = _f64(X), _f64(U)
    out = loss_and_grads(Xd, Ud * Xd, Ud, weight, bandwidth, mults, need_dX=False, block=block)
    out["dU"] = out["dY"] * Xd + out["dU_penalty"]
    return out


# --------------------------------------------------------------------------------------
# torch_two_sample.MMDStatistic restatement  (parity unpinned, see module docstring)
# --------------------------------------------------------------------------------------

def mmd_statistic(s1, s2, alphas: Sequence[float], eps: float = 1e-5):
    """Unbiased MMD^2 with k = sum_a exp(-a * (d2 + eps)); returns (mmd, K[n,n]).

    torch_two_sample/statistics_diff.py::MMDStatistic.__call__ with util.pdist(norm=2):
    a00 = 1/(n1(n1-1)), a11 = 1/(n2(n2-1)), a01 = -1/(n1 n2);
    mmd = 2 a01 sum(K12) + a00 (sum(K11) - tr K11) + a11 (sum(K22) - tr K22).
    """
    A, B = _f64(s1), _f64(s2)
    n1, n2 = A.shape[0], B.shape[0]
    Z = torch.cat([A, B], 0)
    na = (Z * Z).sum(1)
    d2 = eps + (na[:, None] + na[None, :] - 2.0 * (Z @ Z.T)).abs()
    K = torch.zeros_like(d2)
    for a in alphas:
        K += torch.exp(-float(a) * d2)
    a00 = 1.0 / (n1 * (n1 - 1))
    a11 = 1.0 / (n2 * (n2 - 1))
    a01 = -1.0 / (n1 * n2)
    k1, k2, k12 = K[:n1, :n1], K[n1:, n1:], K[:n1, n1:]
    mmd = 2 * a01 * float(k12.sum()) + a00 * float(k1.sum() - torch.trace(k1)) \
        + a11 * float(k2.sum() - torch.trace(k2))
    return mmd, K


def permutation_statistics(K, n1: int, n2: int, perms: np.ndarray) -> np.ndarray:
    """Statistic under each relabeling ``perms[p]`` (a permutation of range(n1+n2);
    the first n1 entries form sample 1).  permutation_test.pyx::permutation_test_mat."""
    K = _f64(K).numpy()
    a00 = 1.0 / (n1 * (n1 - 1))
    a11 = 1.0 / (n2 * (n2 - 1))
    a01 = -1.0 / (n1 * n2)
    out = np.empty(len(perms))
    diag = np.diag(K)
    for p, perm in enumerate(perms):
        i1, i2 = perm[:n1], perm[n1:]
        k11 = K[np.ix_(i1, i1)].sum() - diag[i1].sum()
        k22 = K[np.ix_(i2, i2)].sum() - diag[i2].sum()
        k12 = K[np.ix_(i1, i2)].sum()
        out[p] = 2 * a01 * k12 + a00 * k11 + a11 * k22
    return out


def permutation_pval(K, n1: int, n2: int, perms: np.ndarray) -> float:
    """Fraction of relabelings whose statistic is >= the observed one (identity labels)."""
    K = _f64(K)
    obs = permutation_statistics(K, n1, n2, np.arange(n1 + n2)[None, :])[0]
    st = permutation_statistics(K, n1, n2, perms)
    return float((st >= obs).sum()) / float(len(perms))


# --------------------------------------------------------------------------------------
# Timed CPU baseline (bench.py cpu_baseline / --impl reference): fp32, multithreaded
# --------------------------------------------------------------------------------------

def cpu_blockwise_fwd_bwd(X: torch.Tensor, U: torch.Tensor, bandwidth: float,
                          row_begin: int, row_end: int, block: int = 2048, weight: float = 0.0):
    """fp32 blockwise port of one masked fwd+bwd evaluation restricted to gradient rows
    [row_begin, row_end) of Y (and the matching 1/share of the S_xx block), so a bounded
    SAMPLE of the N=65536 workload can be timed on host cores.  Same math as
    ``mmd2_terms``; torch ops use all host threads.  Returns (partial sums, dU rows)."""
    X = X.float()
    Y = (U * X).float()
    N = X.shape[0]
    mult = torch.tensor(bandwidth_multipliers(), dtype=torch.float32)
    s = scales(bandwidth, mult.tolist())
    sums = [0.0, 0.0, 0.0]
    cx, cy = 1.0 / N, -1.0 / N
    nX = (X * X).sum(1)
    nY = (Y * Y).sum(1)
    dU = torch.empty(row_end - row_begin, X.shape[1])
    for i0 in range(row_begin, row_end, block):
        i1 = min(i0 + block, row_end)
        Yi, Xi = Y[i0:i1], X[i0:i1]
        O = torch.zeros_like(Yi)
        r = torch.zeros(i1 - i0)
        for j0 in range(0, N, block):
            j1 = min(j0 + block, N)
            for (B, nB, cj, which) in ((X[j0:j1], nX[j0:j1], cx, 1), (Y[j0:j1], nY[j0:j1], cy, 2)):
                d2 = (nY[i0:i1, None] + nB[None, :] - 2.0 * (Yi @ B.T)).clamp_min_(0)
                k = torch.zeros_like(d2)
                w = torch.zeros_like(d2)
                for sm in s:
                    e = torch.exp(d2 * (-1.0 / sm))
                    k += e
                    w.add_(e, alpha=1.0 / sm)
                sums[which] += float(k.sum())
                w *= cj
                O += w @ B
                r += w.sum(1)
            # the same share of the X x X block (value only, no gradient)
            d2 = (nX[i0:i1, None] + nX[None, j0:j1] - 2.0 * (Xi @ X[j0:j1].T)).clamp_min_(0)
            k = torch.zeros_like(d2)
            for sm in s:
                k += torch.exp(d2 * (-1.0 / sm))
            sums[0] += float(k.sum())
        dU[i0 - row_begin:i1 - row_begin] = 4.0 * cy * (O - r[:, None] * Yi) * Xi
    if weight != 0:
        # weight * mean_d(1 - max_i U[i,d]) and its gradient (Mmd_loss_constrained.py:50), over the whole mask
        top = torch.topk(U.float(), 1, dim=0)
        sums.append(float(weight) * float((1.0 - top.values).mean()))
        rows = top.indices.reshape(-1)
        inside = (rows >= row_begin) & (rows < row_end)
        cols = torch.arange(U.shape[1])[inside]
        dU[rows[inside] - row_begin, cols] -= float(weight) / U.shape[1]
    return sums, dU


def bench_sample_fp64(X: torch.Tensor, U: torch.Tensor, bandwidth: float, row_begin: int, row_end: int,
                      xx_tile: int, tile: int = 128, block: int = 4096):
    """fp64 truth for a SAMPLE of one masked evaluation at sizes where the full oracle would take hours
    (N = 65536, D = 1024): the gradient rows [row_begin, row_end) of Y = U*X against all 2N columns (their S_xy, S_yy
    share and dL/dU rows, Mmd_loss_constrained.py:46-50 + autograd) and the share of S_xx that the X row tile
    ``xx_tile`` owns when the symmetric X x X block is walked by its upper triangle: the diagonal tile once, every
    tile to its right twice.  Also the fp32 port's own dU error on the same rows (the reference-precision yard-stick).
    """
    N = X.shape[0]
    Xd = X.double()
    Yd = (U.double() * Xd)
    s = scales(bandwidth, bandwidth_multipliers())
    cx, cy = 1.0 / N, -1.0 / N
    nX = (Xd * Xd).sum(1)
    nY = (Yd * Yd).sum(1)
    Yi = Yd[row_begin:row_end]
    O = torch.zeros_like(Yi)
    r = torch.zeros(row_end - row_begin, dtype=torch.float64)
    sxy = syy = 0.0
    for j0 in range(0, N, block):
        j1 = min(j0 + block, N)
        for (B, nB, cj, is_x) in ((Xd[j0:j1], nX[j0:j1], cx, True), (Yd[j0:j1], nY[j0:j1], cy, False)):
            d2 = (nY[row_begin:row_end, None] + nB[None, :] - 2.0 * (Yi @ B.T)).clamp_min_(0)
            k = torch.zeros_like(d2)
            w = torch.zeros_like(d2)
            for sm in s:
                e = torch.exp(-d2 / sm)
                k += e
                w += e / sm
            if is_x:
                sxy += float(k.sum())
            else:
                syy += float(k.sum())
            w *= cj
            O += w @ B
            r += w.sum(1)
    dU = 4.0 * cy * (O - r[:, None] * Yi) * Xd[row_begin:row_end]
    t0, t1 = xx_tile * tile, min((xx_tile + 1) * tile, N)
    Xi = Xd[t0:t1]
    sxx = 0.0
    for j0 in range(t0, N, block):
        j1 = min(j0 + block, N)
        d2 = (nX[t0:t1, None] + nX[None, j0:j1] - 2.0 * (Xi @ Xd[j0:j1].T)).clamp_min_(0)
        k = torch.zeros_like(d2)
        for sm in s:
            k += torch.exp(-d2 / sm)
        wgt = torch.where(torch.arange(j0, j1) < t1, 1.0, 2.0).double()
        sxx += float((k * wgt[None, :]).sum())
    _, dU32 = cpu_blockwise_fwd_bwd(X, U, bandwidth, row_begin, row_end)
    ref32 = float((dU32.double() - dU).norm() / dU.norm())
    return dict(sxy=sxy, syy=syy, dU=dU, sxx_tile=sxx, ref_fp32_dU=ref32)
