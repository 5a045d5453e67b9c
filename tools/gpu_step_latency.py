"""Training-step latency of VGAN_no_kl.fit at the reference's own configuration (config 1: 1000 x 20, batch 500)
and an MNIST-shaped one (batch 4096, D = 784): eager launches vs CUDA-graph replay.  Usage: python tools/gpu_step_latency.py"""
import contextlib, io, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vgan_b200 import VGAN_no_kl

def run(X, batch, epochs, graph):
    VGAN_no_kl.cuda_graph = graph
    m = VGAN_no_kl(batch_size=batch, epochs=epochs, lr=0.01, seed=777)
    with contextlib.redirect_stdout(io.StringIO()):
        m.fit(X[: batch * 2])                      # warm caches / JIT
        torch.cuda.synchronize(); t0 = time.perf_counter()
        m.fit(X)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
    steps = epochs * (len(X) // batch)
    return dt / steps * 1e3, m.train_history["generator_loss"][-1]

rs = np.random.RandomState(0)
for n, d, batch, epochs in ((1000, 20, 500, 200), (16384, 784, 4096, 50)):
    X = rs.randn(n, d).astype(np.float32)
    for graph in (False, True):
        ms, last = run(X, batch, epochs, graph)
        print(f"N={batch:5d} D={d:4d} {'graph' if graph else 'eager'}: {ms:7.3f} ms/step   final epoch loss {last:.6f}", flush=True)
