"""Three evaluations of every given shape, for an ncu launch list:
python tools/gpu_shape_launches.py N D precision [N D precision ...]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import make_inputs
from vgan_b200 import MMDLossConstrained, RBF
from vgan_b200.dist import sharded_masked_step
a = sys.argv[1:]
for k in range(0, len(a), 3):
    N, D, prec = int(a[k]), int(a[k + 1]), a[k + 2]
    X, U = make_inputs(N, D, seed=1)
    X, U = X.cuda(), U.cuda()
    f = MMDLossConstrained(10.0, kernel=RBF(), precision=prec)
    for _ in range(3):
        sharded_masked_step(f, X, U, time_k1=False)
    torch.cuda.synchronize()
    del X, U, f
