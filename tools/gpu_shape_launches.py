"""One evaluation of a given shape for an ncu launch list: python tools/gpu_shape_launches.py N D precision"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import make_inputs
from vgan_b200 import MMDLossConstrained, RBF
from vgan_b200.dist import sharded_masked_step
N, D, prec = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
X, U = make_inputs(N, D, seed=1)
X, U = X.cuda(), U.cuda()
f = MMDLossConstrained(10.0, kernel=RBF(), precision=prec)
for _ in range(3):
    sharded_masked_step(f, X, U, time_k1=False)
torch.cuda.synchronize()
