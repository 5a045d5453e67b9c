"""Subspace distribution on the device.

``unique_subspaces(u)`` = ``np.unique(u, axis=0, return_counts=True)`` of the boolean mask matrix that
``generate_subspaces`` returns (/root/reference/src/vgan.py:355-382, 639-659): rows in lexicographic
order (False < True, column 0 most significant) and their multiplicities, computed by
``vgan_b200_subspace_dist`` without moving the masks to the host."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib


def unique_subspaces(u: torch.Tensor, threshold: float | None = None):
    """u: bool/uint8 masks [n, D], or the generator's float output [n, D] with ``threshold``
    (default 1/D, vgan.py:369).  Returns (unique_rows bool [m, D], counts int64 [m]) on u's device."""
    if u.dim() != 2:
        raise RuntimeError("unique_subspaces: expected a 2-D mask matrix")
    if u.device.type != "cuda":
        raise RuntimeError("vgan_b200.unique_subspaces needs a CUDA (sm_100) tensor; there is no CPU path")
    n, D = u.shape
    L = _lib.lib()
    if u.dtype in (torch.bool, torch.uint8):
        src, is_bool, thr = u.to(torch.uint8).contiguous(), 1, 0.0
    else:
        src, is_bool = u.detach().to(torch.float32).contiguous(), 0
        thr = float(1.0 / D if threshold is None else threshold)
    nbytes = C.c_size_t()
    _lib.check(L.vgan_b200_subspace_scratch_bytes(n, D, C.byref(nbytes)), "subspace_scratch_bytes")
    scratch = torch.empty(nbytes.value + 16, dtype=torch.uint8, device=u.device)
    uniq = torch.empty((n, D), dtype=torch.uint8, device=u.device)
    counts = torch.empty(n, dtype=torch.int64, device=u.device)
    n_unique = torch.zeros(1, dtype=torch.int32, device=u.device)
    with torch.cuda.device(u.device):
        ctx = _lib.ctx(u.device.index if u.device.index is not None else torch.cuda.current_device())
        _lib.check(L.vgan_b200_subspace_dist(ctx, src.data_ptr(), is_bool, n, D, thr, scratch.data_ptr(), uniq.data_ptr(),
                                             counts.data_ptr(), n_unique.data_ptr(),
                                             torch.cuda.current_stream(u.device).cuda_stream), "subspace_dist")
    m = int(n_unique.item())                # the one host sync: the result shape
    return uniq[:m].bool(), counts[:m]
