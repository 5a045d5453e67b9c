"""Warm per-step times of one evaluation (stats / scales / pack / fused fwd+bwd / finalize), each of the five C-ABI
calls bracketed by CUDA events inside a loop of evaluations (inputs and workspace stay in L2 where they fit — what a
training loop sees; the ncu launch lists are cold-cache and serialised):
python tools/gpu_step_breakdown.py N D precision [N D precision ...]"""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import make_inputs
from vgan_b200 import _lib
from vgan_b200.loss import _workspace, _ctx, _stream, _ptr, make_flags, default_precision, RBF

L = _lib.lib()
a = sys.argv[1:]
for k in range(0, len(a), 3):
    N, D, prec = int(a[k]), int(a[k + 1]), a[k + 2]
    X, U = make_inputs(N, D, seed=1)
    X, U = X.cuda(), U.cuda()
    dev = X.device
    kernel = RBF()
    mults = kernel._mults
    flags = make_flags(prec, D, mults, True, False)
    ws = _workspace(N, N, D, D, mults, flags, dev)
    ctx, st, lay = _ctx(dev), _stream(dev), ws.layout_ref
    bw = torch.zeros(1, device=dev)
    dU = torch.empty((N, D), dtype=torch.float32, device=dev)
    reps = 30 if N * N * D < 1e12 else 5
    names = ["stats", "scales", "pack", "fwd_bwd", "finalize"]
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(6)] for _ in range(reps)]
    for it in range(reps + 3):
        e = ev[max(it - 3, 0)]
        e[0].record()
        _lib.check(L.vgan_b200_stats(ctx, lay, ws.ptr, _ptr(X), _ptr(U), None, 0, N, st), "stats"); e[1].record()
        _lib.check(L.vgan_b200_scales(ctx, lay, ws.ptr, _ptr(bw), 0, ws.mults_c, 10.0, st), "scales"); e[2].record()
        _lib.check(L.vgan_b200_pack(ctx, lay, ws.ptr, _ptr(X), _ptr(U), None, 0, N, st), "pack"); e[3].record()
        _lib.check(L.vgan_b200_mmd2_fwd_bwd(ctx, lay, ws.ptr, _ptr(X), _ptr(U), None, 0, N, 0, 1, _ptr(dU), _ptr(X),
                                            None, st), "mmd2_fwd_bwd"); e[4].record()
        _lib.check(L.vgan_b200_finalize(ctx, lay, ws.ptr, 10.0, _ptr(dU), 0, N, 1, None, st), "finalize"); e[5].record()
    torch.cuda.synchronize()
    us = {n: 1e3 * sum(e[i].elapsed_time(e[i + 1]) for e in ev) / reps for i, n in enumerate(names)}
    us["total"] = 1e3 * sum(e[0].elapsed_time(e[5]) for e in ev) / reps
    # the same five calls replayed as one CUDA graph: the evaluation without host launch gaps
    def five():
        _lib.check(L.vgan_b200_stats(ctx, lay, ws.ptr, _ptr(X), _ptr(U), None, 0, N, st2), "stats")
        _lib.check(L.vgan_b200_scales(ctx, lay, ws.ptr, _ptr(bw), 0, ws.mults_c, 10.0, st2), "scales")
        _lib.check(L.vgan_b200_pack(ctx, lay, ws.ptr, _ptr(X), _ptr(U), None, 0, N, st2), "pack")
        _lib.check(L.vgan_b200_mmd2_fwd_bwd(ctx, lay, ws.ptr, _ptr(X), _ptr(U), None, 0, N, 0, 1, _ptr(dU), _ptr(X),
                                            None, st2), "mmd2_fwd_bwd")
        _lib.check(L.vgan_b200_finalize(ctx, lay, ws.ptr, 10.0, _ptr(dU), 0, N, 1, None, st2), "finalize")
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        st2 = side.cuda_stream
        five()
        side.synchronize()
        with torch.cuda.graph(g, stream=side):
            five()
    torch.cuda.synchronize()
    for _ in range(3):
        g.replay()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(reps):
        g.replay()
    t1.record()
    torch.cuda.synchronize()
    us["graph_total"] = 1e3 * t0.elapsed_time(t1) / reps
    print(json.dumps({"N": N, "D": D, "precision": prec, "us": {k: round(v, 1) for k, v in us.items()}}), flush=True)
    del X, U, dU
