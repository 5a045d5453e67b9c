import sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
from vgan_b200 import MMDLossConstrained, RBF
N, D, prec = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
rs = np.random.RandomState(0)
X = torch.tensor(rs.randn(N, D).astype(np.float32), device="cuda"); U = torch.tensor(rs.rand(N, D).astype(np.float32), device="cuda", requires_grad=True)
f = MMDLossConstrained(0.0, kernel=RBF(), precision=prec)
try:
    loss = f.forward_masked(X, U); loss.backward(); torch.cuda.synchronize()
    print(N, D, prec, "ok", float(loss))
except Exception as e:
    print(N, D, prec, "FAIL", str(e)[:60])
