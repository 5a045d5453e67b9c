"""Config 2/3/5 sweep on one B200: fused fwd+bwd evaluations/s of vgan_b200 (bf16, and the two fp32-class paths:
bf16x3 = split operands on the tensor cores, simt = CUDA cores)
next to torch-eager running the reference formula on the same GPU (where its [5,2N,2N] tensor
fits).  Prints one JSON line per (N, D)."""
import json, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vgan_b200 import MMDLossConstrained, RBF
from oracle import torch_eager

def _peak_tf():
    try:
        return float(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                 "MEASURED_PEAKS.json")))["bf16_tflops_sustained"])
    except Exception:
        return 1399.3

PEAK_TF = _peak_tf()

def timeit(fn, reps):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps

def main():
    shapes = [(4096, 784), (4096, 49), (16384, 32), (16384, 128), (16384, 512), (4096, 16), (4096, 64), (4096, 256),
              (4096, 1024), (4096, 4096), (16384, 1024), (65536, 64), (65536, 256), (65536, 1024), (131072, 128),
              (262144, 16), (262144, 64)]
    for N, D in shapes:
        g = torch.Generator(device="cuda").manual_seed(0)
        X = torch.randn(N, D, device="cuda", generator=g)
        p = torch.softmax(torch.randn(N, D, device="cuda", generator=g), 1)
        U0 = torch.where(p < 1.0 / D, p, torch.ones_like(p))
        row = {"N": N, "D": D}
        for prec in ("bf16", "bf16x3", "simt"):
            if prec == "simt" and N * N * D > 16384 * 16384 * 512: continue
            if prec == "bf16x3" and N > 131072: continue
            loss_fn = MMDLossConstrained(10.0, kernel=RBF(), precision=prec)
            U = U0.clone().requires_grad_()
            def step():
                U.grad = None
                loss_fn.forward_masked(X, U).backward()
            reps = 20 if N * N * D < 1e12 else 5
            ms = timeit(step, reps)
            row[f"{prec}_ms"] = round(ms, 4)
            row[f"{prec}_evals_per_s"] = round(1e3 / ms, 2)
            if prec == "bf16":
                row["bf16_tflops_alg"] = round(8.0 * N * N * D / (ms * 1e-3) / 1e12, 1)
                # fractions of the two rooflines of SURVEY.md 8(d): tensor pipe (measured sustained bf16 peak) and MUFU
                # (2 N^2 ex2 per evaluation against 16 ex2/clk/SM x 148 SMs x 1.965 GHz); the larger one binds
                row["tensor_frac"] = round(row["bf16_tflops_alg"] / PEAK_TF, 3)
                row["mufu_frac"] = round(2.0 * N * N / (ms * 1e-3) / (16 * 148 * 1.965e9), 3)
        if 5 * (2 * N) ** 2 * 4 * 4 < 150e9:      # [5,2N,2N] fp32 + autograd copies must fit
            try:
                ref = torch_eager.MMDLossConstrained(10.0)
                U = U0.clone().requires_grad_()
                def rstep():
                    U.grad = None
                    ref(X, U * X, U).backward()
                ms = timeit(rstep, 3)
                row["torch_eager_ms"] = round(ms, 3)
                row["speedup_vs_torch_eager_bf16"] = round(ms / row["bf16_ms"], 1)
                del ref
            except torch.OutOfMemoryError:
                row["torch_eager_ms"] = "oom"
            torch.cuda.empty_cache()
        print(json.dumps(row), flush=True)

if __name__ == "__main__":
    main()
