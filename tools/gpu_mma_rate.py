"""tcgen05.mma rate on resident operands, alone and with concurrent TMA fills of the same shared memory."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vgan_b200 import _lib
L = _lib.lib_debug()      # probes live in the debug build only
CTX = _lib.ctx(0, L)
rows = 32768
Z = torch.randn(rows, 1024, device="cuda").bfloat16()
grid = 148
out = torch.zeros(2 * grid, dtype=torch.int64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
iters = 4000
for n in (128, 256):
    for mn in (0, 1):
        for tw in (0, 1, 2, 4, 8):
            _lib.check(L.vgan_b200_umma_rate_probe(CTX, n, iters, mn, Z.data_ptr(), rows, tw, grid, out.data_ptr(), st), "rate")
            torch.cuda.synchronize()
            o = out.view(grid, 2).double()
            cyc = o[:, 0].mean().item(); boxes = o[:, 1].mean().item()
            print(f"N={n:3d} B {'MN' if mn else 'K '}-major, {tw} TMA warps: {cyc/(iters*4):6.1f} clk/MMA (ideal {n//2}), "
                  f"operand reads {(4096 + n*32)/ (cyc/(iters*4)):6.1f} B/clk, TMA fills {boxes*16384/cyc:6.1f} B/clk")
