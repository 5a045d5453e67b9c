import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vgan_b200 import _lib
L = _lib.lib()
out = torch.zeros(1, dtype=torch.int64, device="cuda")
for n in (64, 128, 192, 256):
    for mn in (0, 1):
        for iters in (2000,):
            _lib.check(L.vgan_b200_umma_rate_probe(n, iters, mn, out.data_ptr(), torch.cuda.current_stream().cuda_stream), "rate")
            torch.cuda.synchronize()
            c = int(out[0])
            print(f"N={n:3d} B {'MN' if mn else 'K '}-major: {c/(iters*4):7.1f} cycles per MMA (ideal {n/2:.0f})")
