"""world_size-2 gloo tests (CPU) of the N>1 host logic: row shards + S_xx tile partition + the single
all-reduce reproduce the 1-rank result; the three exchanges of the data-parallel form work on plain tensors.  Per-shard partials come from the oracle (test
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


def _exchange_worker(rank, world, port, ret):
    """The three exchanges of the data-parallel form (dist.py) on plain tensors: statistics sum + column-max keys,
    all-gather of packed rows, result all-reduce."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from vgan_b200.dist import exchange_statistics, gather_rows, reduce_results, shard_rows
    rs = np.random.RandomState(1)
    N, D = 512, 6
    n = N // world
    X, U = rs.randn(N, D), rs.rand(N, D) - 0.5            # negative mask values too: the key order must survive
    b, e = rank * n, (rank + 1) * n
    ok = shard_rows(N, rank, world) == (b, e)
    # (1) per-rank statistics [sq_x, sq_y | colsum x | colsum y] and packed (value, ~row) keys
    Xl, Yl = X[b:e], (U * X)[b:e]
    stats = torch.tensor(np.concatenate([[np.sum(Xl ** 2), np.sum(Yl ** 2)], Xl.sum(0), Yl.sum(0)]))

    def key(v, row):                                         # prologue.cu: float_key << 32 | (0xffffffff - row)
        bits = np.float32(v).view(np.uint32).astype(np.uint64)
        fk = np.where(bits & 0x80000000, ~bits & 0xffffffff, bits | 0x80000000).astype(np.uint64)
        return ((fk << np.uint64(32)) | np.uint64(0xffffffff - row)).astype(np.uint64)
    keys = np.zeros(D, dtype=np.uint64)
    for r in range(b, e):
        keys = np.maximum(keys, key(U[r], r))
    cmax = torch.tensor(keys.view(np.int64))
    exchange_statistics(stats, cmax, None)
    want_stats = np.concatenate([[np.sum(X ** 2), np.sum((U * X) ** 2)], X.sum(0), (U * X).sum(0)])
    ok = ok and np.allclose(stats.numpy(), want_stats, rtol=1e-12)
    rows = 0xffffffff - (cmax.numpy().view(np.uint64) & np.uint64(0xffffffff)).astype(np.int64)
    ok = ok and np.array_equal(rows, np.argmax(U.astype(np.float32), axis=0))
    # (2) packed operand rows: every rank fills its slot, the gather completes the array
    Z = torch.zeros(N, D)
    Z[b:e] = torch.tensor(X[b:e], dtype=torch.float32)
    gather_rows(Z, rank, world, None, in_place=False)
    ok = ok and torch.equal(Z, torch.tensor(X, dtype=torch.float32))
    # (3) [S_xx, S_xy, S_yy, spare | colsum]
    red = torch.arange(4 + D, dtype=torch.float64) * (rank + 1)
    reduce_results(red, None)
    ok = ok and torch.equal(red, torch.arange(4 + D, dtype=torch.float64) * sum(range(1, world + 1)))
    ret[rank] = bool(ok)
    dist.destroy_process_group()


def test_data_parallel_exchanges():
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_exchange_worker, args=(2, 29541, ret), nprocs=2, join=True)
    assert ret[0] and ret[1]
