"""``MMDStatistic`` with the call convention of torch_two_sample (josipd/torch-two-sample,
unpinned and absent in the reference environment) as V-GAN uses it in ``check_if_myopic``
(/root/reference/src/vgan.py:419-428, 696-705):

    mmd = MMDStatistic(n_1, n_2)
    value, K = mmd(sample_1, sample_2, alphas=[a], ret_matrix=True)
    p = mmd.pval(K, n_permutations=1000)

Unbiased (diagonal-removed) MMD^2 with k = sum_a exp(-a (d2 + 1e-5)).  The kernel matrix comes
from the CUDA ``mmd2_matrix`` kernel; the permutation test evaluates all relabelings on the device
(``vgan_b200_permutation_stats``, csrc/permtest.cu: 8 relabelings per CTA share every K_ij read) instead of
the upstream Cython O(n_perm n^2) host loop.
Parity for this class is UNPINNED (no upstream source or vectors available): it is checked
against the restatement in oracle/mmd_oracle.py only."""
from __future__ import annotations

import torch

from . import _lib
from .loss import _as_input, _ctx, _ptr, _stream


class MMDStatistic:
    def __init__(self, n_1, n_2):
        self.n_1 = n_1
        self.n_2 = n_2
        self.a00 = 1. / (n_1 * (n_1 - 1))
        self.a11 = 1. / (n_2 * (n_2 - 1))
        self.a01 = - 1. / (n_1 * n_2)

    def _statistic(self, K, member_1):
        """member_1: uint8 [P, n] rows marking sample 1 with ones; returns the [P] statistics (fp64) from the
        device permutation kernel (csrc/permtest.cu) — no torch matmul on this path."""
        import ctypes as C
        K = K.contiguous()
        member_1 = member_1.contiguous()
        n = K.shape[0]
        L = _lib.lib()
        nbytes = C.c_size_t()
        _lib.check(L.vgan_b200_permutation_scratch_bytes(n, C.byref(nbytes)), "permutation_scratch_bytes")
        scratch = torch.empty(nbytes.value, dtype=torch.uint8, device=K.device)
        stats = torch.empty(member_1.shape[0], dtype=torch.float64, device=K.device)
        with torch.cuda.device(K.device):
            _lib.check(L.vgan_b200_permutation_stats(_ctx(K.device), _ptr(K), n, self.n_1, _ptr(member_1),
                                                     member_1.shape[0], _ptr(scratch), _ptr(stats), _stream(K.device)),
                       "permutation_stats")
        return stats

    def __call__(self, sample_1, sample_2, alphas, ret_matrix=False):
        if sample_1.device.type != "cuda":
            raise RuntimeError("vgan_b200.MMDStatistic needs CUDA tensors on an sm_100 device")
        dev = sample_1.device
        Z = torch.cat([_as_input(sample_1, dev), _as_input(sample_2, dev)], 0).contiguous()
        n, D = Z.shape
        if n != self.n_1 + self.n_2:
            raise RuntimeError("MMDStatistic: sample sizes do not match n_1, n_2")
        inv = torch.tensor([float(a) for a in alphas], dtype=torch.float32, device=dev)
        if inv.numel() < 1 or inv.numel() > _lib.MAX_SCALES:
            raise RuntimeError(f"MMDStatistic: between 1 and {_lib.MAX_SCALES} alphas are supported")
        K = torch.empty((n, n), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().vgan_b200_mmd2_matrix(_ctx(dev), _ptr(Z), n, D, _ptr(inv), inv.numel(), 1e-5, _ptr(K),
                                                        _stream(dev)), "mmd2_matrix")
        member = torch.zeros(1, n, dtype=torch.uint8, device=dev)
        member[0, :self.n_1] = 1
        mmd = self._statistic(K, member)[0].float()
        if ret_matrix:
            return mmd, K
        return mmd

    def pval(self, distances, n_permutations=1000):
        """Fraction of random relabelings whose statistic is >= the observed one."""
        if distances.device.type != "cuda":
            raise RuntimeError("vgan_b200.MMDStatistic needs CUDA tensors on an sm_100 device")
        K = distances.float()
        n = K.shape[0]
        dev = K.device
        member = torch.zeros(n_permutations + 1, n, dtype=torch.uint8, device=dev)
        member[0, :self.n_1] = 1
        order = torch.rand(n_permutations, n, device=dev).argsort(1)       # the relabelings (RNG stays in torch)
        member[1:].scatter_(1, order[:, :self.n_1], 1)
        st = self._statistic(K, member)
        return float((st[1:] >= st[0]).double().mean().item())
