"""world_size-2 gloo test (CPU) of the N>1 host logic: row shards + S_xx tile partition + the single
all-reduce reproduce the 1-rank result.  Per-shard partials come from the oracle (test
infrastructure) since no GPU exists here; the GPU kernels use the same shard arguments."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import mmd_oracle as orc


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from vgan_b200.dist import combine_shards, shard_rows
    rs = np.random.RandomState(0)
    N, D = 384, 12
    X, U = rs.randn(N, D), rs.rand(N, D)
    Y = U * X
    bw = orc.bandwidth_direct(np.vstack([X, Y]))
    b, e = shard_rows(N, rank, world)
    # this rank: Y rows [b,e) against all columns, S_xx row tiles t = rank (mod world)
    Xt, Yt = torch.tensor(X), torch.tensor(Y)
    _, sxy, syy, _, dY = orc.mmd2_terms(Xt, Yt, bw, need_dX=False)     # full, then slice this shard
    K = orc.rbf_matrix(torch.cat([Xt, Yt]), bw)
    tiles = [t for t in range((N + 127) // 128) if t % world == rank]
    sxx = sum(float(K[t * 128:min((t + 1) * 128, N), :N].sum()) for t in tiles)
    sxy_p = float(K[N + b:N + e, :N].sum())
    syy_p = float(K[N + b:N + e, N:].sum())
    dU = (dY * Xt)[b:e]
    pen, _ = orc.colmax_penalty(U, 10.0)
    loss, s, colsum = combine_shards(torch.tensor([sxx, sxy_p, syy_p]), dU.sum(0), torch.tensor(pen), N, N)
    ref = orc.masked_loss_and_grad(X, U, 10.0, bandwidth=bw)
    ok = abs(float(loss) - ref["loss"]) < 1e-6 and abs(float(s[1]) - sxy) < 1e-6 * sxy and \
        torch.allclose(colsum, (ref["dY"] * Xt).sum(0), atol=1e-9)
    ret[rank] = bool(ok)
    dist.destroy_process_group()


def test_two_rank_shards_reduce_to_single_rank_result():
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, 29533, ret), nprocs=2, join=True)
    assert ret[0] and ret[1]


def _gather_worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from vgan_b200.dist import gather_batch, shard_rows
    rs = np.random.RandomState(1)
    N, D = 512, 6
    X, U = torch.tensor(rs.randn(N, D)), torch.tensor(rs.rand(N, D))
    n = N // world
    Xg, Ug, off = gather_batch(X[rank * n:(rank + 1) * n], U[rank * n:(rank + 1) * n])
    ok = torch.equal(Xg, X) and torch.equal(Ug, U) and off == rank * n and shard_rows(N, rank, world) == (off, off + n)
    try:
        gather_batch(X[:100], U[:100])
        ok = False
    except RuntimeError:
        pass
    ret[rank] = bool(ok)
    dist.destroy_process_group()


def test_data_parallel_gather_lines_up_with_row_shards():
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_gather_worker, args=(2, 29541, ret), nprocs=2, join=True)
    assert ret[0] and ret[1]
