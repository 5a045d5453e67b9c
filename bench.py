#!/usr/bin/env python
"""bench.py — fused MMD^2 forward+backward evaluations/sec (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W          # our arm (CUDA, sm_100a)
    python bench.py --impl reference ...                    # reference CPU path (oracle port)

Workload (configs[3], the one the metric is quoted on): one V-GAN large-batch step,
N = 65536 samples, D = 1024 features, X ~ N(0,1), U = upper_softmax(randn) (Generator.py:18-22),
weight = 10 (VGAN_no_kl, vgan.py:571), 5 bandwidths.  One "step" = one evaluation:
X, U -> loss, dL/dU [N, D], colsum(dL/dU) [D]  through vgan_b200.dist.sharded_masked_step — the same C-ABI calls
MMDLossConstrained.forward_masked(...).backward() makes, without the autograd wrapper (no kernel differs).
N > 1: one process per GPU (torchrun); gradient rows and the S_xx tiles sharded across ranks.  `value` feeds
resident, replicated inputs (one NCCL all-reduce of [S_xx,S_xy,S_yy | colsum(dU)] per step) or, when faster, the
data-parallel form (own-row prologue, all-gather of the packed bf16 operand); `e2e` copies only the rank's own rows
from pinned host memory and always uses the data-parallel form.  The default run adds a `secondary` list (rank 0,
N = 1): BASELINE.json configs[1], configs[2], the fp32-class path at the bench shape and torch-eager on the same GPU.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_DEFAULT, D_DEFAULT, WEIGHT = 65536, 1024, 10.0
WORKLOAD = "configs[3]: large-batch MMD2 step N={N}, D={D}, 5 bandwidths, weight=10"


def make_inputs(N, D, seed=0):
    g = torch.Generator().manual_seed(seed)
    X = torch.randn(N, D, generator=g)
    h = torch.randn(N, D, generator=g)
    p = torch.softmax(h, 1)
    U = torch.where(p < 1.0 / D, p, torch.ones_like(p))
    return X, U


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            m = json.load(f)
        return float(m["bf16_tflops_sustained"]), "measured(sustained)"
    except Exception:
        return 1400.0, "fallback"


class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons = index, False, [], set()
        self.max_mhz = None

    def run(self):
        while not self.stop_flag:
            self._once()
            time.sleep(0.2)

    def _once(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        if True:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                parts = [s.strip() for s in out.strip().split(",")]
                self.samples.append(float(parts[0]))
                self.max_mhz = float(parts[1])
                for nm, v in zip(names, parts[2:]):
                    if v.lower().startswith("active"):
                        self.reasons.add(nm)
            except Exception:
                pass

    def result(self):
        if not self.samples:
            self.stop_flag = False
            self._once()                       # timed region shorter than one sampling period: sample right after it
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


def host_threads():
    """All host threads for the CPU legs (torchrun exports OMP_NUM_THREADS=1 to its workers)."""
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    if torch.get_num_threads() < n:
        torch.set_num_threads(n)
    return torch.get_num_threads()


def run_reference(args):
    """Reference arm: the CPU restatement of the reference path (oracle port; the reference module
    itself cannot hold its [5,2N,2N] tensor at this shape and does not travel to the GPU box),
    all host threads, each step = a bounded row sample of the same workload (penalty term included)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import mmd_oracle as orc
    N, D = args.n, args.d
    X, U = make_inputs(N, D)
    bw = 2.0 * D   # value is irrelevant for timing; ~ E|zi-zj|^2
    cores = host_threads()
    rows = args.ref_rows
    if rows <= 0:
        # size the per-step sample so that the whole --steps/--warmup run takes about two minutes on this host:
        # a 2048-row probe gives the rate (the cost is linear in the rows), then whole 2048-row blocks up to N/2
        t0 = time.perf_counter()
        orc.cpu_blockwise_fwd_bwd(X, U, bw, 0, 2048, weight=WEIGHT)
        per_row = (time.perf_counter() - t0) / 2048
        rows = int(120.0 / (args.steps + args.warmup) / per_row) // 2048 * 2048
        rows = max(2048, min(rows, N // 2))
    times = []
    for it in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        orc.cpu_blockwise_fwd_bwd(X, U, bw, 0, rows, weight=WEIGHT)
        dt = time.perf_counter() - t0
        if it >= args.warmup:
            times.append(dt)
    t = float(np.mean(times))
    evals = (rows / N) / t
    sample = (f"gradient rows [0,{rows}) of N={N} against all 2N columns + the matching share of the XxX block + the "
              f"column-max penalty; {t:.2f} s of CPU work per step, value and ms_per_step scaled by N/rows")
    line = {"impl": "reference", "metric": "mmd2_fwd_bwd_evals_per_sec", "value": evals, "unit": "evals/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t * N / rows,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": {"workload": WORKLOAD.format(N=N, D=D)},
            "sample_rows": rows, "sample_fraction": rows / N, "measured_ms_per_step": 1e3 * t,
            "cpu_baseline": {"value": evals, "unit": "evals/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": evals, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def time_steps(fn, steps, warmup):
    """ms per call of fn() on the current stream, CUDA events, after warm-up."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / steps


def micro_peaks():
    """profiles/peaks_r2.json: MUFU / FP32 / epilogue-mix / L2 / tcgen05 ceilings measured on this pool's B200 by
    tools/peaks (committed; the file is named in the output)."""
    try:
        with open(os.path.join(ROOT, "profiles", "peaks_r2.json")) as f:
            return json.load(f)
    except Exception:
        return None


def secondary_configs(dev, peak_tf):
    """BASELINE.json configs[1], configs[2], the fp32-class path at the bench shape and the torch-eager arm, each timed
    for a few steps on one GPU: ms per evaluation, the roofline that binds it and the fraction of it.
    tensor: 8 N^2 D flops / measured sustained bf16 peak.  epilogue: 2.5 N^2 kernel pairs per evaluation (N x 2N
    gradient rows + the upper triangle of X x X) / the measured pairs/s of the ladder epilogue's instruction mix
    (2 ex2 + 17 packed fp32 per 2 pairs: MUFU and FP32 issue together, profiles/peaks_r2.json; that is the round-1
    epilogue and stays the yardstick — the kernel now issues 14 to 16 packed operations per column pair)."""
    from vgan_b200 import MMDLossConstrained, RBF
    from vgan_b200.dist import sharded_masked_step
    pk = micro_peaks()
    epi_peak = pk["ladder_epilogue_mix"]["pairs_per_s"] if pk else None
    mufu_peak = pk["mufu_ex2"]["per_s"] if pk else None
    out = []

    def ours(name, N, D, prec, steps=8, warmup=3):
        X, U = make_inputs(N, D, seed=1)
        X, U = X.to(dev), U.to(dev)
        f = MMDLossConstrained(WEIGHT, kernel=RBF(), precision=prec)
        ms = time_steps(lambda: sharded_masked_step(f, X, U, 0, 1, time_k1=False), steps, warmup)
        graph_ms = None
        if N <= 16384:
            # the same evaluation replayed as a CUDA graph (what VGAN.fit does): no host launch gaps — at these sizes the
            # per-call number above depends on the host as much as on the GPU
            try:
                side = torch.cuda.Stream()
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    sharded_masked_step(f, X, U, 0, 1, time_k1=False)
                    side.synchronize()
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=side):
                        keep = sharded_masked_step(f, X, U, 0, 1, time_k1=False)
                torch.cuda.current_stream().wait_stream(side)
                graph_ms = time_steps(g.replay, 3 * steps, warmup)
                del g, keep
            except Exception:
                graph_ms = None
        t = ms * 1e-3
        tensor_frac = 8.0 * N * N * D / t / 1e12 / peak_tf
        rec = {"config": name, "N": N, "D": D, "precision": prec, "ms": ms, "evals_per_s": 1e3 / ms,
               "tensor_frac": tensor_frac}
        if graph_ms is not None:
            rec["graph_replay_ms"] = graph_ms
        if epi_peak:
            rec["epilogue_frac"] = 2.5 * N * N / t / epi_peak
            rec["mufu_frac"] = 2.0 * N * N / t / mufu_peak
        fr = {"tensor": tensor_frac, "epilogue(mufu+fp32 issue)": rec.get("epilogue_frac", 0.0)}
        rec["bound"] = max(fr, key=fr.get)
        rec["frac"] = fr[rec["bound"]]
        out.append(rec)
        del X, U
        return ms

    ours("configs[1]: MNIST-shaped, batch 4096, D=784", 4096, 784, "bf16", steps=20)
    ours("configs[1]: its VGAN encodings, D=49 (default precision: fp32-class)", 4096, 49, "fp32", steps=20)
    for D in (32, 128, 512):
        for prec in ("bf16", "fp32"):
            ours(f"configs[2]: ADBench-shaped N=16384, D={D}, {'bf16-operand' if prec == 'bf16' else 'fp32-class (split operands)'}",
                 16384, D, prec)
    ours("configs[3] shape on the fp32-class path (split bf16 operands, 1e-4)", N_DEFAULT, D_DEFAULT, "fp32", steps=3, warmup=1)
    # the reference formula (torch eager, cuBLAS SGEMM + elementwise kernels + autograd) on the same GPU: the on-box
    # "kernel to beat" (SURVEY.md 2.3); its [5,2N,2N] fp32 tensor is 20 GiB at N = 16384
    try:
        from oracle import torch_eager
        N, D = 16384, 512
        X, U0 = make_inputs(N, D, seed=1)
        X, U0 = X.to(dev), U0.to(dev)
        ref = torch_eager.MMDLossConstrained(WEIGHT, kernel=torch_eager.RBF(device=dev), device=dev)

        def eager_step():
            U = U0.clone().requires_grad_()
            ref(X, U * X, U).backward()
        ms = time_steps(eager_step, 3, 1)
        mine = next(r["ms"] for r in out if r["N"] == N and r["D"] == D and r["precision"] == "fp32")
        out.append({"config": "torch-eager reference formula (oracle/torch_eager.py) on the same B200, fp32", "N": N, "D": D,
                    "precision": "fp32 eager", "ms": ms, "evals_per_s": 1e3 / ms,
                    "tensor_frac": 8.0 * N * N * D / (ms * 1e-3) / 1e12 / peak_tf, "bound": "hbm (unfused)", "frac": None,
                    "speedup_of_fp32_class_path": ms / mine})
        del X, U0, ref
        torch.cuda.empty_cache()
    except Exception as exc:       # e.g. out of memory on a smaller part: reported, not fatal
        out.append({"config": "torch-eager reference formula", "error": str(exc)[:200]})
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=N_DEFAULT)
    ap.add_argument("--d", type=int, default=D_DEFAULT)
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--cpu-rows", type=int, default=16384,
                    help="gradient rows of the cpu_baseline sample of our arm (16384 rows ~ 10 s on 16 threads)")
    ap.add_argument("--ref-rows", type=int, default=0,
                    help="--impl reference: gradient rows per step (0 = sized so the whole run takes ~2 minutes)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-secondary", action="store_true")
    ap.add_argument("--feed", default="auto", choices=["auto", "replicated", "gathered"],
                    help="N>1, device-timed value: resident replicated inputs, or own rows + all-gather of the packed operand")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    # stdout carries exactly ONE line (the JSON): whatever libraries print there (NCCL's version banner at N > 1)
    # goes to stderr instead
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch.distributed as dist
    from vgan_b200 import MMDLossConstrained, RBF, _lib
    from vgan_b200.dist import data_parallel_masked_step, sharded_masked_step

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    N, D = args.n, args.d
    Xh, Uh = make_inputs(N, D)
    Xh, Uh = Xh.pin_memory(), Uh.pin_memory()
    X, U = Xh.to(dev), Uh.to(dev)
    loss_fn = MMDLossConstrained(WEIGHT, kernel=RBF(), precision=args.precision)
    can_gather = world > 1 and N % (128 * world) == 0
    n_loc = N // world if can_gather else N
    Xl, Ul = (X[rank * n_loc:(rank + 1) * n_loc], U[rank * n_loc:(rank + 1) * n_loc]) if can_gather else (X, U)

    def step_replicated():
        return sharded_masked_step(loss_fn, X, U, rank, world)

    def step_gathered():
        return data_parallel_masked_step(loss_fn, Xl, Ul, time_k1=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(v):
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    # ---- which feed for the device-timed value (N > 1): both are probed for a few steps, the faster one is timed
    feed = "replicated"
    if world > 1 and can_gather:
        if args.feed == "gathered":
            feed = "gathered"
        elif args.feed == "auto":
            probe = {}
            for name, fn in (("replicated", step_replicated), ("gathered", step_gathered)):
                for _ in range(2):
                    fn()
                barrier()
                t0 = time.perf_counter()
                for _ in range(3):
                    fn()
                barrier()
                probe[name] = allmax(time.perf_counter() - t0)
            feed = min(probe, key=probe.get)
    step = step_gathered if feed == "gathered" else step_replicated

    # ---- device-resident timing
    for _ in range(args.warmup):
        out = step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k_ms = []
    launches0 = _lib.launch_count(local)
    ev0.record()
    for _ in range(args.steps):
        out = step()
        k_ms.append(out["k1_events"])
    ev1.record()
    barrier()
    launches = _lib.launch_count(local) - launches0
    sampler.stop_flag = True
    ms = allmax(ev0.elapsed_time(ev1) / args.steps)
    k1_ms = allmax(float(np.mean([a.elapsed_time(b) for a, b in k_ms])))
    loss_val = float(out["loss"])

    # ---- end to end: pinned host inputs -> H2D -> step -> D2H of loss + colsum(dU)
    e2e = None
    if not args.no_e2e:
        # Two device input sets: the H2D copy of step i+1 runs on a copy stream under the kernels of step i
        # (every step still pays for its own H2D and its D2H read-back inside the timed region).
        # N > 1: every rank copies only ITS rows; the packed bf16 operand rows are all-gathered over NVLink inside the
        # step (dist.data_parallel_masked_step).  If N is not a multiple of 128 x ranks every rank copies everything.
        res_h = torch.empty(1 + D, dtype=torch.float64).pin_memory()
        copy_st = torch.cuda.Stream(device=dev)
        main_st = torch.cuda.current_stream()
        shard = can_gather
        r0 = rank * n_loc if shard else 0
        Xsrc, Usrc = Xh[r0:r0 + n_loc], Uh[r0:r0 + n_loc]
        bufs = [(torch.empty((n_loc, D), device=dev), torch.empty((n_loc, D), device=dev)) for _ in range(2)]
        ready = [torch.cuda.Event(), torch.cuda.Event()]
        done = [torch.cuda.Event(), torch.cuda.Event()]
        res_d = torch.empty(1 + D, dtype=torch.float64, device=dev)

        def h2d(i):
            Xd, Ud = bufs[i & 1]
            with torch.cuda.stream(copy_st):
                copy_st.wait_event(done[i & 1])        # the step that last read this buffer set has finished
                Xd.copy_(Xsrc, non_blocking=True)
                Ud.copy_(Usrc, non_blocking=True)
                ready[i & 1].record(copy_st)

        def e2e_step(i):
            h2d(i + 1)                                 # next step's inputs, overlapped with this step
            Xd, Ud = bufs[i & 1]
            main_st.wait_event(ready[i & 1])
            if shard:
                o = data_parallel_masked_step(loss_fn, Xd, Ud)
            else:
                o = sharded_masked_step(loss_fn, Xd, Ud, rank, world, time_k1=False)
            done[i & 1].record(main_st)
            res_d[:1].copy_(o["loss"].reshape(1))
            res_d[1:].copy_(o["dU_colsum"])
            res_h.copy_(res_d, non_blocking=True)
            main_st.synchronize()                      # the user reads the loss every step (vgan.py:620)

        for e in done:
            e.record(main_st)
        h2d(0)
        for i in range(2):
            e2e_step(i)
        barrier()
        t0 = time.perf_counter()
        for i in range(2, 2 + args.steps):
            e2e_step(i)
        barrier()
        e_ms = allmax((time.perf_counter() - t0) * 1e3 / args.steps)
        del bufs
        e2e = {"value": 1e3 / e_ms, "unit": "evals/s", "h2d_bytes_per_step": 2 * n_loc * D * 4 * world,
               "d2h_bytes_per_step": (1 + D) * 8 * world, "ms_per_step": e_ms,
               "overlap": "H2D of step i+1 on a copy stream under the kernels of step i; loss + colsum(dU) read every step",
               "feed": (f"every rank copies its own N/{world} rows; packed bf16 operand rows all-gathered over NVLink inside the step"
                        if shard else "every rank copies the full batch")}

    if rank == 0:
        peak, peak_src = peaks()
        flops = 8.0 * N * N * D / world            # algorithmic flops of this rank's launch
        achieved = flops / (k1_ms * 1e-3) / 1e12
        traffic, traffic_src = None, None
        try:      # DRAM read+write bytes of one gradient-kernel launch from the committed ncu --set full capture
            with open(os.path.join(ROOT, "profiles", "r2_pair_traffic.json")) as f:
                tj = json.load(f)
            if (N, D, world) == (tj["N"], tj["D"], 1):
                traffic, traffic_src = tj["dram_bytes_per_launch"], tj["source"]
        except Exception:
            pass
        line = {"metric": "mmd2_fwd_bwd_evals_per_sec", "value": 1e3 / ms, "unit": "evals/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "f32",
                "data": "synthetic",
                "config": {"workload": WORKLOAD.format(N=N, D=D),
                           "l2": "inputs (X,U 512 MiB; bf16 operand 256 MiB) larger than L2",
                           "parallelism": f"row-tile shards x{world}" + ("" if world == 1 else
                                          (", own-row prologue + all-gather of the packed bf16 operand" if feed == "gathered"
                                           else ", resident replicated inputs, one all-reduce of [sums | colsum dU]")),
                           "loss": loss_val},
                "gpu_launches": launches,
                "gpu_launches_note": "kernels launched by libvgan_b200.so on rank 0 inside the timed region, counted by the "
                                     "library (vgan_b200_launch_count); NCCL's own kernels are not included",
                "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                             "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                             "peak_source": peak_src,
                             "kernel": "mmd2_tc_pair_kernel (gradient units) with mmd2_xx_kernel (S_xx sums) overlapped",
                             "kernel_ms": k1_ms,
                             "algorithmic": "8*N^2*D flops per evaluation (SURVEY.md 8(d))"},
                "clocks": sampler.result()}
        if e2e:
            line["e2e"] = e2e
        if world == 1 and not args.no_secondary and (N, D) == (N_DEFAULT, D_DEFAULT):
            del X, U
            torch.cuda.empty_cache()
            line["secondary"] = secondary_configs(dev, peak)
            line["secondary_peaks"] = "profiles/peaks_r2.json (tools/peaks) and MEASURED_PEAKS.json bf16_tflops_sustained"
        if not args.no_cpu_baseline and world == 1:     # reported at N=1 only (rank 0 has the host to itself there)
            from oracle import mmd_oracle as orc
            host_threads()
            rows = args.cpu_rows
            t0 = time.perf_counter()
            orc.cpu_blockwise_fwd_bwd(Xh, Uh, 2.0 * D, 0, rows, weight=WEIGHT)
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": (rows / N) / dt, "unit": "evals/s", "cores": torch.get_num_threads(),
                                    "kind": "port",
                                    "sample": f"gradient rows [0,{rows}) of N={N} vs all 2N columns + matching "
                                              f"share of XxX + penalty, {dt:.1f}s of CPU work, scaled by N/rows"}
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
