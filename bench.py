#!/usr/bin/env python
"""bench.py — fused MMD^2 forward+backward evaluations/sec (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W          # our arm (CUDA, sm_100a)
    python bench.py --impl reference ...                    # reference CPU path (oracle port)

Workload (configs[3], the one the metric is quoted on): one V-GAN large-batch step,
N = 65536 samples, D = 1024 features, X ~ N(0,1), U = upper_softmax(randn) (Generator.py:18-22),
weight = 10 (VGAN_no_kl, vgan.py:571), 5 bandwidths.  One "step" = one evaluation:
X, U -> loss, dL/dU  through MMDLossConstrained.forward_masked(...).backward().
N > 1: one process per GPU (torchrun), inputs replicated, gradient rows and the S_xx tiles
sharded across ranks, one NCCL all-reduce of [S_xx,S_xy,S_yy | colsum(dU)] per step.
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
KERNELS_PER_STEP = 10  # colstats, mean, pack, scales, setup, K1 (pair kernel), K1s (S_xx sums), colmax, penalty, finalize


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
    all host threads, each step = a bounded row sample of the same workload."""
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
        orc.cpu_blockwise_fwd_bwd(X, U, bw, 0, 2048)
        per_row = (time.perf_counter() - t0) / 2048
        rows = int(120.0 / (args.steps + args.warmup) / per_row) // 2048 * 2048
        rows = max(2048, min(rows, N // 2))
    times = []
    for it in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        orc.cpu_blockwise_fwd_bwd(X, U, bw, 0, rows)
        dt = time.perf_counter() - t0
        if it >= args.warmup:
            times.append(dt)
    t = float(np.mean(times))
    evals = (rows / N) / t
    sample = f"gradient rows [0,{rows}) of N={N} against all 2N columns + the matching share of the XxX block"
    line = {"impl": "reference", "metric": "mmd2_fwd_bwd_evals_per_sec", "value": evals, "unit": "evals/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t * N / rows,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": {"workload": f"configs[3]: large-batch MMD2 step N={N}, D={D}, 5 bandwidths"},
            "cpu_baseline": {"value": evals, "unit": "evals/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": evals, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


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
    ap.add_argument("--dp-gather", action="store_true",
                    help="N>1: every rank holds only its rows of X and U; the step all-gathers them first (NCCL)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    # stdout carries exactly ONE line (the JSON): whatever libraries print there (NCCL's version banner at N > 1)
    # goes to stderr instead
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch.distributed as dist
    from vgan_b200 import MMDLossConstrained, RBF
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

    def step(Xd, Ud):
        if args.dp_gather and world > 1:
            n = N // world
            return data_parallel_masked_step(loss_fn, Xd[rank * n:(rank + 1) * n], Ud[rank * n:(rank + 1) * n])
        return sharded_masked_step(loss_fn, Xd, Ud, rank, world)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing
    for _ in range(args.warmup):
        out = step(X, U)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k_ms = []
    ev0.record()
    for _ in range(args.steps):
        out = step(X, U)
        k_ms.append(out["k1_events"])
    ev1.record()
    barrier()
    sampler.stop_flag = True
    ms = ev0.elapsed_time(ev1) / args.steps
    k1_ms = float(np.mean([a.elapsed_time(b) for a, b in k_ms]))
    t = torch.tensor([ms, k1_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, k1_ms = float(t[0]), float(t[1])
    loss_val = float(out["loss"])

    # ---- end to end: pinned host inputs -> H2D -> step -> D2H of loss + colsum(dU)
    e2e = None
    if not args.no_e2e:
        # Two device input sets: the H2D copy of step i+1 runs on a copy stream under the kernels of step i
        # (every step still pays for its own H2D and its D2H read-back inside the timed region).
        # N > 1 has two ways to feed the host batch: "replicated" = every rank copies the whole batch (N x the H2D
        # bytes, no exchange) or "sharded" = every rank copies only ITS rows and the shards are all-gathered over NVLink
        # inside the step (dist.gather_batch).  Both are probed for 2 steps; the faster one is timed and named.
        res_h = torch.empty(1 + D, dtype=torch.float32).pin_memory()
        copy_st = torch.cuda.Stream(device=dev)
        main_st = torch.cuda.current_stream()

        def run_e2e(shard, steps):
            n_loc = N // world if shard else N
            r0 = rank * n_loc if shard else 0
            Xsrc, Usrc = Xh[r0:r0 + n_loc], Uh[r0:r0 + n_loc]
            bufs = [(torch.empty((n_loc, D), device=dev), torch.empty((n_loc, D), device=dev)) for _ in range(2)]
            ready = [torch.cuda.Event(), torch.cuda.Event()]
            done = [torch.cuda.Event(), torch.cuda.Event()]

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
                o = data_parallel_masked_step(loss_fn, Xd, Ud) if shard else step(Xd, Ud)
                done[i & 1].record(main_st)
                res_h[:1].copy_(o["loss"].reshape(1), non_blocking=True)
                res_h[1:].copy_(o["dU_colsum"], non_blocking=True)
                main_st.synchronize()                      # the user reads the loss every step (vgan.py:620)

            for e in done:
                e.record(main_st)
            h2d(0)
            for i in range(2):
                e2e_step(i)
            barrier()
            t0 = time.perf_counter()
            for i in range(2, 2 + steps):
                e2e_step(i)
            barrier()
            e_ms = (time.perf_counter() - t0) * 1e3 / steps
            te = torch.tensor([e_ms], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(te, op=dist.ReduceOp.MAX)
            del bufs
            return float(te[0])

        shard = False
        if world > 1 and N % (128 * world) == 0:
            shard = run_e2e(True, 2) < run_e2e(False, 2)       # same decision on every rank (all-reduced times)
        e_ms = run_e2e(shard, args.steps)
        e2e = {"value": 1e3 / e_ms, "unit": "evals/s", "h2d_bytes_per_step": 2 * N * D * 4,
               "d2h_bytes_per_step": (1 + D) * 4,
               "overlap": "H2D of step i+1 on a copy stream under the kernels of step i; loss + colsum(dU) read every step",
               "feed": ("sharded: every rank copies its N/%d rows, shards all-gathered over NVLink inside the step" % world)
                       if shard else "replicated: every rank copies the full batch"}

    if rank == 0:
        peak, peak_src = peaks()
        flops = 8.0 * N * N * D / world            # algorithmic flops of this rank's launch
        achieved = flops / (k1_ms * 1e-3) / 1e12
        line = {"metric": "mmd2_fwd_bwd_evals_per_sec", "value": 1e3 / ms, "unit": "evals/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "f32",
                "data": "synthetic",
                "config": {"workload": f"configs[3]: large-batch MMD2 step N={N}, D={D}, 5 bandwidths, weight=10",
                           "l2": "inputs (X,U 512 MiB; bf16 operand 256 MiB) larger than L2",
                           "parallelism": f"row-tile shards x{world}" + (", inputs all-gathered per step" if args.dp_gather else ""),
                           "loss": loss_val},
                "gpu_launches": KERNELS_PER_STEP * args.steps,
                "roofline": {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                             "frac": achieved / peak,
                             # DRAM read+write bytes of one pair-kernel launch, ncu --set full (profiles/r1h_pair_k1_summary.txt)
                             "traffic": (11.22e9 / world) if (N, D) == (N_DEFAULT, D_DEFAULT) else None, "peak_source": peak_src,
                             "kernel": "mmd2_tc_pair_kernel (gradient units) with mmd2_xx_kernel (S_xx sums) overlapped",
                             "kernel_ms": k1_ms,
                             "algorithmic": "8*N^2*D flops per evaluation (SURVEY.md 8(d))"},
                "clocks": sampler.result()}
        if e2e:
            line["e2e"] = e2e
        if not args.no_cpu_baseline and world == 1:     # reported at N=1 only (rank 0 has the host to itself there)
            from oracle import mmd_oracle as orc
            host_threads()
            rows = args.cpu_rows
            Xc, Uc = Xh[:], Uh[:]
            t0 = time.perf_counter()
            orc.cpu_blockwise_fwd_bwd(Xc, Uc, 2.0 * D, 0, rows)
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": (rows / N) / dt, "unit": "evals/s", "cores": torch.get_num_threads(),
                                    "kind": "port",
                                    "sample": f"gradient rows [0,{rows}) of N={N} vs all 2N columns + matching "
                                              f"share of XxX, {dt:.1f}s of CPU work, scaled by N/rows"}
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
