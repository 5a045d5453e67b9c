"""One MMD^2 evaluation across the GPUs of a node, one process per GPU.

The pairwise matrix is split by row tiles: rank g computes the gradient rows R_g = [g*N/G, (g+1)*N/G) of Y = U*X
against ALL columns (complete dL/dU rows, no reduction needed), the partial sums S_xy, S_yy of those rows and the
S_xx tiles t = g (mod G).  Two forms, same kernels:

* ``sharded_masked_step``: X and U are resident on every rank (replicated generator).  Every rank runs the HBM-bound
  prologue over all rows (cheaper than moving the packed operand at these sizes) and K1 over its own rows; the ONE
  exchange is an all-reduce of [S_xx, S_xy, S_yy, - | colsum_i dL/dU[i,:]] (4 + D_p doubles, straight out of the
  kernels' accumulator block) over NCCL/NVLink.
* ``data_parallel_masked_step``: every rank holds only ITS rows of the batch and of the generator output (its own
  noise).  The prologue runs on the rank's own rows; exchanged are (1) the statistics block (all-reduce sum: Σ|x|^2,
  Σ|y|^2, column sums -> bandwidth, centring mean, common mode) and the column-max keys of U (all-reduce max),
  (2) the packed bf16 operand rows and scaled norms (all-gather: half the bytes of the fp32 inputs), (3) the result
  block as above.  Reference step being distributed: /root/reference/src/vgan.py:597-621.

The exchange helpers take plain tensors and work on any backend (NCCL on the GPUs, gloo in the CPU tests)."""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import _lib
from .loss import _ctx, _ptr, _stream, _workspace, default_precision, make_flags, mmd2_eval

_I64_MIN = -2 ** 63


def shard_rows(N: int, rank: int, world: int, tile: int = 128):
    """Contiguous, tile-aligned row range of rank `rank` (the last ranks may get fewer rows, or none)."""
    tiles = (N + tile - 1) // tile
    per = (tiles + world - 1) // world
    b = min(rank * per * tile, N)
    e = min((rank + 1) * per * tile, N)
    return b, e


# ------------------------------------------------------------------ the three exchanges
def exchange_statistics(stats: torch.Tensor, cmax, group=None):
    """(1) in place: sum of the per-rank statistics [Σ|x|^2, Σ|y|^2 | colsum X | colsum Y] (float64) and maximum of the
    packed (value, ~row) column-max keys of U.  The keys are unsigned 64-bit on the device; flipping the top bit makes
    their order the signed order the collective's MAX uses."""
    dist.all_reduce(stats, group=group)
    if cmax is not None:
        flip = torch.tensor(_I64_MIN, dtype=torch.int64, device=cmax.device)
        keys = torch.bitwise_xor(cmax, flip)
        dist.all_reduce(keys, op=dist.ReduceOp.MAX, group=group)
        cmax.copy_(torch.bitwise_xor(keys, flip))


def gather_rows(full: torch.Tensor, rank: int, world: int, group=None, in_place: bool = True):
    """(2) all-gather of equal row chunks into ``full`` [world * n, ...]; this rank's chunk is already in place
    (NCCL's in-place form: the send buffer is the rank's slot of the receive buffer)."""
    n = full.shape[0] // world
    mine = full[rank * n:(rank + 1) * n]
    dist.all_gather_into_tensor(full, mine if in_place else mine.clone(), group=group)


def reduce_results(reduced: torch.Tensor, group=None):
    """(3) in place: all-reduce of [S_xx, S_xy, S_yy, spare | colsum dL/dU] (float64) — the one exchange per step
    BASELINE.json's north star names (the scalars plus the D-vector)."""
    dist.all_reduce(reduced, group=group)


def combine_shards(sums, dU_colsum, penalty, Nx: int, Ny: int, group=None):
    """Host-side form of (3) on separate tensors (kept for callers that hold partial results outside a workspace):
    returns (loss, sums, colsum)."""
    packed = torch.cat([sums[:3].double(), dU_colsum.double()])
    dist.all_reduce(packed, group=group)
    s = packed[:3]
    loss = (s[0] / (Nx * Nx) - 2.0 * s[1] / (Nx * Ny) + s[2] / (Ny * Ny)).float() + penalty
    return loss, s, packed[3:]


# ------------------------------------------------------------------ replicated inputs
def sharded_masked_step(loss_fn, X, U, rank: int = 0, world: int = 1, group=None, time_k1: bool = True):
    """loss(X, U*X, U) and dL/dU rows of this rank; returns dict(loss, dU, rows, dU_colsum, k1_events, sums).

    With world == 1 this is exactly ``loss_fn.forward_masked(X, U)`` + backward, without autograd.  A rank whose shard
    is empty (more ranks than row tiles) still runs its S_xx tiles and takes part in the all-reduce."""
    N, D = X.shape
    dev = X.device
    kernel = loss_fn.kernel
    bw, frozen = kernel._bw_state(dev)
    prec = loss_fn.precision or default_precision(D)
    weight = float(loss_fn.weight)
    b, e = shard_rows(N, rank, world)
    ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) if time_k1 else None
    if world == 1:
        out = mmd2_eval(X, None, U, weight=weight, bw_dev=bw, frozen=frozen, mults=kernel._mults, precision=prec,
                        masked=True, center=loss_fn.center, k1_events=ev, xx_nparts=1, row_begin=0, row_end=N)
        kernel._freeze()
        return dict(loss=out["loss"][0], dU=out["dY"], rows=(0, N), dU_colsum=out["colsum"], k1_events=ev,
                    sums=out["sums"])
    L = _lib.lib()
    use_pen = weight != 0
    flags = make_flags(prec, D, kernel._mults, loss_fn.center, False)
    with torch.cuda.device(dev):
        ws = _workspace(N, N, D, D if use_pen else 0, kernel._mults, flags, dev)
        ctx, st, lay = _ctx(dev), _stream(dev), ws.layout_ref
        dU = torch.empty((e - b, D), dtype=torch.float32, device=dev) if e > b else None
        _lib.check(L.vgan_b200_stats(ctx, lay, ws.ptr, _ptr(X), _ptr(U), None, 0, N, st), "stats")
        _lib.check(L.vgan_b200_scales(ctx, lay, ws.ptr, _ptr(bw), 1 if frozen else 0, ws.mults_c, weight, st), "scales")
        _lib.check(L.vgan_b200_pack(ctx, lay, ws.ptr, _ptr(X), _ptr(U), None, 0, N, st), "pack")
        if ev:
            ev[0].record()
        _lib.check(L.vgan_b200_mmd2_fwd_bwd(ctx, lay, ws.ptr, _ptr(X), _ptr(U), None, b, e, rank, world, _ptr(dU),
                                            _ptr(X[b:]) if dU is not None else None, None, st), "mmd2_fwd_bwd")
        if ev:
            ev[1].record()
        reduce_results(ws.reduced, group)
        _lib.check(L.vgan_b200_finalize(ctx, lay, ws.ptr, weight, _ptr(dU), b, e - b, 1, None, st), "finalize")
    kernel._freeze()
    return dict(loss=ws.loss[0], dU=dU, rows=(b, e), dU_colsum=ws.colsum, k1_events=ev, sums=ws.sums)


# ------------------------------------------------------------------ data parallel: every rank holds its own rows
def data_parallel_masked_step(loss_fn, X_local, U_local, group=None, time_k1: bool = False):
    """loss(X, U*X, U) over the global batch whose rows are spread over the ranks (equal shards, a multiple of 128
    rows each); returns this rank's rows of dL/dU (``dU`` lines up with ``U_local``), the global loss and the global
    colsum(dL/dU).  Prologue on the rank's own rows, then the three exchanges of the module docstring."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    n, D = X_local.shape
    if world == 1:
        return sharded_masked_step(loss_fn, X_local, U_local, 0, 1, group, time_k1)
    if n % 128:
        # the same test on every rank (equal shards): all ranks raise together, nobody is left in a collective
        raise RuntimeError("data_parallel_masked_step: every rank's shard must be a multiple of 128 rows")
    N = n * world
    b, e = rank * n, (rank + 1) * n
    dev = X_local.device
    kernel = loss_fn.kernel
    bw, frozen = kernel._bw_state(dev)
    prec = loss_fn.precision or default_precision(D)
    weight = float(loss_fn.weight)
    use_pen = weight != 0
    flags = make_flags(prec, D, kernel._mults, loss_fn.center, False)
    L = _lib.lib()
    ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) if time_k1 else None
    with torch.cuda.device(dev):
        ws = _workspace(N, N, D, D if use_pen else 0, kernel._mults, flags, dev)
        ctx, st, lay = _ctx(dev), _stream(dev), ws.layout_ref
        # the C ABI indexes rows of the GLOBAL batch: hand it the address global row 0 would have (only rows [b, e) of
        # it are ever touched)
        Xv = X_local.data_ptr() - b * D * 4
        Uv = U_local.data_ptr() - b * D * 4
        dU = torch.empty((n, D), dtype=torch.float32, device=dev)
        _lib.check(L.vgan_b200_stats(ctx, lay, ws.ptr, Xv, Uv, None, b, e, st), "stats")
        exchange_statistics(ws.stats, ws.cmax, group)
        _lib.check(L.vgan_b200_scales(ctx, lay, ws.ptr, _ptr(bw), 1 if frozen else 0, ws.mults_c, weight, st), "scales")
        _lib.check(L.vgan_b200_pack(ctx, lay, ws.ptr, Xv, Uv, None, b, e, st), "pack")
        for z in ws.operand:                         # X block and Y block of the stacked operand, then the norms
            gather_rows(z[:N], rank, world, group)
            gather_rows(z[N:], rank, world, group)
        gather_rows(ws.norms[:N], rank, world, group)
        gather_rows(ws.norms[N:], rank, world, group)
        if ev:
            ev[0].record()
        _lib.check(L.vgan_b200_mmd2_fwd_bwd(ctx, lay, ws.ptr, Xv, Uv, None, b, e, rank, world, _ptr(dU), _ptr(X_local),
                                            None, st), "mmd2_fwd_bwd")
        if ev:
            ev[1].record()
        reduce_results(ws.reduced, group)
        _lib.check(L.vgan_b200_finalize(ctx, lay, ws.ptr, weight, _ptr(dU), b, n, 1, None, st), "finalize")
    kernel._freeze()
    return dict(loss=ws.loss[0], dU=dU, rows=(b, e), dU_colsum=ws.colsum, k1_events=ev, sums=ws.sums)
