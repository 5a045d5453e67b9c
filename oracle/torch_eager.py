"""Torch-eager restatement of the reference loss module, line by line
(/root/reference/src/models/Mmd_loss_constrained.py:5-50), usable on any torch device.
TEST / MEASUREMENT INFRASTRUCTURE ONLY: it is the "torch-eager on B200" arm of the sweep
 and a checker; the product never imports it."""
import torch
from torch import nn


class RBF(nn.Module):
    def __init__(self, n_kernels=5, mul_factor=2.0, bandwidth=None, device="cuda"):
        super().__init__()
        self.bandwidth_multipliers = (mul_factor ** (torch.arange(n_kernels) - n_kernels // 2)).to(device)
        self.bandwidth = bandwidth

    def get_bandwidth(self, L2_distances):
        if self.bandwidth is None:
            n_samples = L2_distances.shape[0]
            self.bandwidth = L2_distances.data.sum() / (n_samples ** 2 - n_samples)
        return self.bandwidth

    def forward(self, X):
        L2_distances = torch.cdist(X, X) ** 2
        return torch.exp(-L2_distances[None, ...] /
                         (self.get_bandwidth(L2_distances) * self.bandwidth_multipliers)[:, None, None]).sum(dim=0)


class MMDLossConstrained(nn.Module):
    def __init__(self, weight, kernel=None, device="cuda"):
        super().__init__()
        self.kernel = RBF(device=device) if kernel is None else kernel
        self.weight = weight
        self.device = device

    def forward(self, X, Y, U):
        K = self.kernel(torch.vstack([X, Y]))
        self.bandwidth = self.kernel.bandwidth
        n = X.shape[0]
        XX, XY, YY = K[:n, :n].mean(), K[:n, n:].mean(), K[n:, n:].mean()
        return XX - 2 * XY + YY + self.weight * torch.mean(
            torch.ones(U.shape[1], device=self.device) - torch.topk(U, 1, 0).values)
