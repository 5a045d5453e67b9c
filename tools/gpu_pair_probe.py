import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vgan_b200 import _lib
L = _lib.lib_debug()      # probes live in the debug build only
CTX = _lib.ctx(0, L)
torch.manual_seed(0)
A0 = torch.randn(128, 128, device="cuda").bfloat16().contiguous()
A1 = torch.randn(128, 128, device="cuda").bfloat16().contiguous()
B = torch.randn(128, 128, device="cuda").bfloat16().contiguous()
V = torch.randn(128, 256, device="cuda").bfloat16().contiguous()
S = torch.zeros(2, 128, 128, device="cuda"); O = torch.zeros(2, 128, 256, device="cuda")
_lib.check(L.vgan_b200_pair_probe(CTX, A0.data_ptr(), A1.data_ptr(), B.data_ptr(), V.data_ptr(), S.data_ptr(), O.data_ptr(),
                                  torch.cuda.current_stream().cuda_stream), "pair_probe")
torch.cuda.synchronize()
for h, A in enumerate((A0, A1)):
    Sr = A.float() @ B.float().T
    Or = S[h].bfloat16().float() @ V.float()
    print(f"CTA {h}: S rel-max-err {float((S[h]-Sr).abs().max()/Sr.abs().max()):.3e}   O rel-max-err {float((O[h]-Or).abs().max()/Or.abs().max()):.3e}")
    if float((S[h]-Sr).abs().max()/Sr.abs().max()) > 1e-2:
        alts = {"other A": (A1 if h == 0 else A0).float() @ B.float().T,
                "B halves swapped": A.float() @ torch.cat([B[64:], B[:64]]).float().T}
        for k, v in alts.items(): print("   S vs", k, float((S[h]-v).abs().max()/v.abs().max()))
    if float((O[h]-Or).abs().max()/Or.abs().max()) > 1e-2:
        Vs = torch.cat([V[:, 128:], V[:, :128]], 1)
        print("   O vs V halves swapped", float((O[h]-S[h].bfloat16().float() @ Vs.float()).abs().max()/Or.abs().max()))
