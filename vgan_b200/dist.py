"""Row-tile sharding of one MMD^2 evaluation across the GPUs of a node.

The pairwise matrix is split by row tiles: rank g computes the gradient rows
[g*N/G, (g+1)*N/G) of Y = U*X against ALL columns (complete dL/dU rows, no reduction needed),
the partial sums S_xy, S_yy of those rows and the S_xx tiles t = g (mod G).  The only exchange
is ONE all-reduce per step of [S_xx, S_xy, S_yy | colsum_i dL/dU[i,:]] (3 + D doubles) over
NCCL/NVLink.  Inputs are replicated (the generator MLP is tiny and data-parallel replicas
produce the same U from the same noise)."""
from __future__ import annotations

import torch
import torch.distributed as dist

from .loss import default_precision, mmd2_eval


def shard_rows(N: int, rank: int, world: int, tile: int = 128):
    """Contiguous, tile-aligned row range of rank `rank` (the last rank takes the remainder)."""
    tiles = (N + tile - 1) // tile
    per = (tiles + world - 1) // world
    b = min(rank * per * tile, N)
    e = min((rank + 1) * per * tile, N)
    return b, e


def combine_shards(sums, dU_colsum, penalty, Nx: int, Ny: int, group=None):
    """The one exchange of the sharded path: all-reduce [S_xx, S_xy, S_yy | colsum dU] (float64)
    and finish the loss.  Works on any backend (NCCL on the GPUs, gloo in the CPU tests)."""
    packed = torch.cat([sums[:3].double(), dU_colsum.double()])
    dist.all_reduce(packed, group=group)
    s = packed[:3]
    loss = (s[0] / (Nx * Nx) - 2.0 * s[1] / (Nx * Ny) + s[2] / (Ny * Ny)).float() + penalty
    return loss, s, packed[3:]


def sharded_masked_step(loss_fn, X, U, rank: int = 0, world: int = 1, group=None, time_k1: bool = True):
    """loss(X, U*X, U) and dL/dU rows of this rank; returns dict(loss, dU, rows, dU_colsum, k1_events).

    With world == 1 this is exactly ``loss_fn.forward_masked(X, U)`` + backward, without autograd."""
    N, D = X.shape
    kernel = loss_fn.kernel
    bw, frozen = kernel._bw_state(X.device)
    prec = loss_fn.precision or default_precision(D)
    b, e = shard_rows(N, rank, world)
    ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) if time_k1 else None
    if e <= b:      # more ranks than row tiles: this rank only joins the all-reduce
        raise RuntimeError("sharded_masked_step: fewer 128-row tiles than ranks")
    out = mmd2_eval(X, None, U, weight=float(loss_fn.weight), bw_dev=bw, frozen=frozen, mults=kernel._mults,
                    precision=prec, need_dX=False, need_dY=True, masked=True, penalty_grad=True,
                    row_begin=b, row_end=e, xx_part=rank, xx_nparts=world, center=loss_fn.center,
                    finalize=(world == 1), k1_events=ev)
    kernel._freeze()
    dU = out["dY"]
    colsum = dU.sum(0, dtype=torch.float64)
    if world > 1:
        # the penalty (column max over ALL rows of the replicated U) is identical on every rank
        loss, _, colsum = combine_shards(out["sums"], colsum, out["penalty"][0], N, N, group)
    else:
        loss = out["loss"][0]
    return dict(loss=loss, dU=dU, rows=(b, e), dU_colsum=colsum.float(), k1_events=ev, sums=out["sums"])


def gather_batch(X_local: torch.Tensor, U_local: torch.Tensor, group=None):
    """Data-parallel form of the step's input: every rank holds ITS rows of the batch and of the generator output
    (its own noise); the column operand of the pairwise matrix is the whole batch, so the shards are all-gathered
    (two ``all_gather_into_tensor`` over NCCL/NVLink; gloo in the CPU tests).  Returns (X, U, row offset of this
    rank).  All shards must have the same number of rows, a multiple of 128 unless world == 1."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return X_local, U_local, 0
    rank = dist.get_rank(group)
    n, D = X_local.shape
    if n % 128:
        raise RuntimeError("gather_batch: every rank's shard must be a multiple of 128 rows (row-tile aligned)")
    X = torch.empty((world * n, D), dtype=X_local.dtype, device=X_local.device)
    U = torch.empty((world * n, D), dtype=U_local.dtype, device=U_local.device)
    dist.all_gather_into_tensor(X, X_local.contiguous(), group=group)
    dist.all_gather_into_tensor(U, U_local.contiguous(), group=group)
    return X, U, rank * n


def data_parallel_masked_step(loss_fn, X_local, U_local, group=None):
    """loss(X, U*X, U) over the global batch whose rows are spread over the ranks; returns this rank's rows of dL/dU
    (``dU`` lines up with ``U_local``) — all-gather of the inputs, then the row-tile sharded step above."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    X, U, _ = gather_batch(X_local, U_local, group)
    out = sharded_masked_step(loss_fn, X, U, rank, world, group)
    b, e = out["rows"]
    n = X_local.shape[0]
    if (b, e) != (rank * n, (rank + 1) * n):
        raise RuntimeError("data_parallel_masked_step: shard rows do not line up with the row-tile partition")
    return out
