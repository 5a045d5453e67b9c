"""``torch.ops.vgan_b200.*`` — the C-ABI entry points registered as PyTorch operators ("torch
binding"): functional ops with schemas, CUDA implementations only.  This file is boundary plumbing (schemas over the
same ctypes calls the nn.Module classes make), not a component of its own.  Calling one with CPU tensors fails in the
dispatcher (there is no CPU kernel to fall back to), with CUDA tensors on a non-sm_100 device in the C library
(``VGAN_B200_ERR_DEVICE``).

    loss, dU, sums        = torch.ops.vgan_b200.mmd2_masked_fwd_bwd(X, U, bandwidth, frozen, mults, weight, precision, center)
    loss, dX, dY, dU, sums = torch.ops.vgan_b200.mmd2_fwd_bwd(X, Y, U, bandwidth, frozen, mults, weight, precision, center)
    K                     = torch.ops.vgan_b200.mmd2_matrix(Z, inv_scales, eps)
    stats                 = torch.ops.vgan_b200.permutation_stats(K, n_1, member)
    rows, counts          = torch.ops.vgan_b200.unique_subspaces(U, threshold)

``bandwidth`` is a 1-element float32 device tensor that the op fills on the first call (``frozen=False``: closed form
of Mmd_loss_constrained.py:16-22) and only reads afterwards, which is how ``RBF`` freezes it.  The ``nn.Module``
classes in loss.py are the drop-in surface; these ops are the same calls without the module state."""
from __future__ import annotations

import torch

from . import loss as _loss
from . import statistic as _statistic
from . import subspace as _subspace
from . import _lib

_LIB = torch.library.Library("vgan_b200", "DEF")
_LIB.define("mmd2_masked_fwd_bwd(Tensor X, Tensor U, Tensor(a!) bandwidth, bool frozen, float[] mults, float weight, "
            "str precision, bool center) -> (Tensor, Tensor, Tensor)")
_LIB.define("mmd2_fwd_bwd(Tensor X, Tensor Y, Tensor? U, Tensor(a!) bandwidth, bool frozen, float[] mults, float weight, "
            "str precision, bool center) -> (Tensor, Tensor, Tensor, Tensor, Tensor)")
_LIB.define("mmd2_matrix(Tensor Z, Tensor inv_scales, float eps) -> Tensor")
_LIB.define("permutation_stats(Tensor K, int n_1, Tensor member) -> Tensor")
_LIB.define("unique_subspaces(Tensor U, float threshold) -> (Tensor, Tensor)")


def _f32(t: torch.Tensor) -> torch.Tensor:
    return t.detach().to(torch.float32).contiguous()


def _mmd2_masked_fwd_bwd(X, U, bandwidth, frozen, mults, weight, precision, center):
    """loss(X, U*X, U) (src/vgan.py:615-616) and dL/dU in one fused pass; sums = [S_xx, S_xy, S_yy, spare] (fp64)."""
    out = _loss.mmd2_eval(_f32(X), None, _f32(U), weight=float(weight), bw_dev=bandwidth, frozen=bool(frozen),
                          mults=list(mults), precision=precision, need_dX=False, need_dY=True, masked=True,
                          penalty_grad=True, center=bool(center))
    return out["loss"][0].clone(), out["dY"], out["sums"].clone()


def _mmd2_fwd_bwd(X, Y, U, bandwidth, frozen, mults, weight, precision, center):
    """MMDLossConstrained.forward(X, Y, U) with all gradients (Mmd_loss_constrained.py:42-50 + autograd)."""
    Uc = _f32(U) if (U is not None and weight != 0) else None
    out = _loss.mmd2_eval(_f32(X), _f32(Y), Uc, weight=float(weight), bw_dev=bandwidth, frozen=bool(frozen),
                          mults=list(mults), precision=precision, need_dX=True, need_dY=True, masked=False,
                          penalty_grad=Uc is not None, center=bool(center))
    dU = out["dU_penalty"] if out["dU_penalty"] is not None else torch.zeros(0, device=X.device)
    return out["loss"][0].clone(), out["dX"], out["dY"], dU, out["sums"].clone()


def _mmd2_matrix(Z, inv_scales, eps):
    """K[n,n] = sum_m exp(-(d2 + eps) * inv_scales[m])  (RBF.forward / MMDStatistic(ret_matrix=True))."""
    Zc, inv = _f32(Z), _f32(inv_scales)
    n, D = Zc.shape
    K = torch.empty((n, n), dtype=torch.float32, device=Zc.device)
    with torch.cuda.device(Zc.device):
        _lib.check(_lib.lib().vgan_b200_mmd2_matrix(_loss._ctx(Zc.device), Zc.data_ptr(), n, D, inv.data_ptr(), inv.numel(),
                                                    float(eps), K.data_ptr(),
                                                    torch.cuda.current_stream(Zc.device).cuda_stream), "mmd2_matrix")
    return K


def _permutation_stats(K, n_1, member):
    """Unbiased MMD^2 statistic of every relabeling (rows of ``member``, uint8) of the kernel matrix K."""
    st = _statistic.MMDStatistic(int(n_1), K.shape[0] - int(n_1))
    return st._statistic(K.to(torch.float32), member.to(torch.uint8))


def _unique_subspaces(U, threshold):
    rows, counts = _subspace.unique_subspaces(U, threshold if threshold > 0 else None)
    return rows, counts


for _name, _fn in (("mmd2_masked_fwd_bwd", _mmd2_masked_fwd_bwd), ("mmd2_fwd_bwd", _mmd2_fwd_bwd),
                   ("mmd2_matrix", _mmd2_matrix), ("permutation_stats", _permutation_stats),
                   ("unique_subspaces", _unique_subspaces)):
    _LIB.impl(_name, _fn, "CUDA")

OPS = ("mmd2_masked_fwd_bwd", "mmd2_fwd_bwd", "mmd2_matrix", "permutation_stats", "unique_subspaces")
