"""L2 -> shared-memory TMA throughput vs bytes in flight (K1's operand-ring access patterns)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vgan_b200 import _lib
L = _lib.lib_debug()      # probes live in the debug build only
CTX = _lib.ctx(0, L)
rows, cols = 131072, 1024
Z = torch.randn(rows, cols, device="cuda").bfloat16()
st = torch.cuda.current_stream().cuda_stream
grid = 148
out = torch.zeros(grid, dtype=torch.int64, device="cuda")
for mode in (2, 6):
    for box_rows in (128,):
        for nprod in (1, 2, 4):
            for stages, loads in ((4, 1), (8, 1), (4, 2), (8, 2), (4, 3), (4, 4)):
                iters = 4096 // loads
                box = box_rows * 128
                if stages * loads * box > 200 * 1024 or stages % nprod: continue
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                args = (Z.data_ptr(), rows, cols, stages, loads, box_rows, iters, mode, grid, nprod, out.data_ptr(), st)
                _lib.check(L.vgan_b200_tma_rate_probe(CTX, *args), "tma")
                a.record()
                _lib.check(L.vgan_b200_tma_rate_probe(CTX, *args), "tma")
                b.record(); torch.cuda.synchronize()
                ms = a.elapsed_time(b)
                byts = iters * loads * box
                cyc = out.double().mean().item()
                print(f"mode={mode} box={box//1024:2d}K nprod={nprod} stages={stages:2d} x loads={loads} inflight={stages*loads*box//1024:3d}K: "
                      f"{byts/cyc:6.1f} B/clk/SM  {cyc/iters:6.0f} clk/stage  {grid*byts/ms/1e9:6.2f} TB/s")
