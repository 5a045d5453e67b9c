"""``upper_softmax`` (the generator's last activation, /root/reference/src/models/Generator.py:18-22) as one CUDA
kernel forward and one backward, feeding the mask straight to the loss prologue:

    p = softmax(h, dim=1);  U = p where p < 1/D else 1;   dh = p * (g - sum_f g_f p_f),  g = dU * [p < 1/D]

The row maximum and 1/sum of the forward are kept so that the backward recomputes p from h instead of saving it."""
from __future__ import annotations

import torch

from . import _lib
from .loss import _ctx, _ptr, _stream


class _UpperSoftmaxFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, h):
        dev = h.device
        hc = h.detach().contiguous()
        N, D = hc.shape
        U = torch.empty_like(hc)
        rmax = torch.empty(N, dtype=torch.float32, device=dev)
        rinv = torch.empty(N, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().vgan_b200_upper_softmax(_ctx(dev), _ptr(hc), N, D, _ptr(U), _ptr(rmax), _ptr(rinv),
                                                          _stream(dev)), "upper_softmax")
        ctx.save_for_backward(hc, rmax, rinv)
        return U

    @staticmethod
    def backward(ctx, g):
        hc, rmax, rinv = ctx.saved_tensors
        N, D = hc.shape
        gc = g.detach().to(torch.float32).contiguous()
        dh = torch.empty_like(hc)
        with torch.cuda.device(hc.device):
            _lib.check(_lib.lib().vgan_b200_upper_softmax_bwd(_ctx(hc.device), _ptr(hc), _ptr(rmax), _ptr(rinv), _ptr(gc),
                                                              N, D, _ptr(dh), _stream(hc.device)), "upper_softmax_bwd")
        return dh


def upper_softmax_cuda(h: torch.Tensor) -> torch.Tensor:
    """h: float32 [N, D] on an sm_100 device."""
    if h.device.type != "cuda" or h.dim() != 2 or h.dtype != torch.float32:
        raise RuntimeError("vgan_b200.upper_softmax_cuda needs a 2-D float32 CUDA tensor")
    return _UpperSoftmaxFn.apply(h)
