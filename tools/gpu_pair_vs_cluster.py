"""Fused step at MNIST-shaped widths (D = 784 -> D_p = 896) through the CTA-pair kernel vs the 1-CTA cluster fallback."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vgan_b200 import MMDLossConstrained, RBF
for N, D in ((16384, 784), (32768, 784), (16384, 320)):
    X = torch.randn(N, D, device="cuda"); U = torch.rand(N, D, device="cuda", requires_grad=True)
    for name, env in (("pair", None), ("cluster fallback", "1")):
        if env: os.environ["VGAN_B200_NO_PAIR"] = env
        else: os.environ.pop("VGAN_B200_NO_PAIR", None)
        f = MMDLossConstrained(10.0, kernel=RBF(), precision="bf16")
        for _ in range(3): U.grad = None; f.forward_masked(X, U).backward()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); a.record()
        for _ in range(10): U.grad = None; f.forward_masked(X, U).backward()
        b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 10
        print(f"N={N} D={D} {name:16s}: {ms:7.3f} ms/step  {8.0*N*N*D/ms/1e9:7.1f} TF/s algorithmic", flush=True)
