"""Dense torch evaluation of V-GAN's squared-MMD loss on any torch device — the "torch-eager on the same B200"
arm of tools/sweep.py and a checker.  TEST / MEASUREMENT INFRASTRUCTURE ONLY; nothing under vgan_b200/ imports it.

It materialises what the reference module materialises (the [n, n] distance matrix of the stacked sample and one
[n, n] kernel layer per bandwidth), so it only runs where that fits; autograd gives the backward.  Follows
/root/reference/src/models/Mmd_loss_constrained.py: multipliers 2^(m-2) (:12-13), bandwidth = sum of squared
distances / (n^2 - n) on the first call, then frozen (:16-22), kernel = sum_m exp(-d^2 / (bandwidth * mult_m)) (:24-26),
biased block means and the column-max penalty (:42-50).  Pinned on CPU against tests/golden/*.npz
(tests/test_oracle.py::test_torch_eager_matches_golden).
"""
import torch


def stacked_sq_distances(Z: torch.Tensor) -> torch.Tensor:
    d = torch.cdist(Z, Z)
    return d * d


def kernel_sum(sq: torch.Tensor, bandwidth: torch.Tensor, multipliers: torch.Tensor) -> torch.Tensor:
    scales = (bandwidth * multipliers).reshape(-1, 1, 1)
    return torch.exp(-sq.unsqueeze(0) / scales).sum(0)


def column_max_penalty(U: torch.Tensor, weight: float) -> torch.Tensor:
    top = torch.topk(U, 1, dim=0).values.reshape(-1)
    return weight * (1.0 - top).mean()


class RBF(torch.nn.Module):
    def __init__(self, n_kernels=5, mul_factor=2.0, bandwidth=None, device="cuda"):
        super().__init__()
        exponents = torch.arange(n_kernels) - n_kernels // 2
        self.bandwidth_multipliers = (mul_factor ** exponents).to(device)
        self.bandwidth = bandwidth

    def forward(self, Z):
        sq = stacked_sq_distances(Z)
        if self.bandwidth is None:                       # first call fixes it for good
            n = sq.shape[0]
            self.bandwidth = sq.detach().sum() / (n * n - n)
        return kernel_sum(sq, self.bandwidth, self.bandwidth_multipliers.to(sq.dtype))


class MMDLossConstrained(torch.nn.Module):
    def __init__(self, weight, kernel=None, device="cuda"):
        super().__init__()
        self.kernel = kernel if kernel is not None else RBF(device=device)
        self.weight = weight
        self.device = device

    def forward(self, X, Y, U):
        nx = X.shape[0]
        K = self.kernel(torch.cat([X, Y], dim=0))
        self.bandwidth = self.kernel.bandwidth
        mmd2 = K[:nx, :nx].mean() - 2.0 * K[:nx, nx:].mean() + K[nx:, nx:].mean()
        return mmd2 + column_max_penalty(U, self.weight)
