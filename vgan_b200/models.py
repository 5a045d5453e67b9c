"""The small torch MLPs that stay in torch (north_star (d)): same layer order, widths and
activation as /root/reference/src/models/Generator.py:6-22,58-70 and Detector.py:5-48, so that
a fixed seed draws identical initial parameters."""
import torch
from torch import nn


class upper_softmax(nn.Module):
    """softmax over features; entries >= 1/D become exactly 1 (Generator.py:18-22).  float32 CUDA activations take
    the fused kernels of softmax.py (one launch forward, one backward); everything else (the MLP is plain torch and
    also runs on the CPU) the torch formula."""

    fused = True

    def forward(self, x):
        if self.fused and x.is_cuda and x.dim() == 2 and x.dtype == torch.float32:
            from .softmax import upper_softmax_cuda
            return upper_softmax_cuda(x)
        p = torch.nn.functional.softmax(x, 1)
        thr = 1 / p.shape[1]
        return torch.less(p, thr) * p + torch.greater_equal(p, thr)


def _mlp(widths):
    return nn.Sequential(*[nn.Linear(a, b) for a, b in zip(widths[:-1], widths[1:])])


class Generator_big(nn.Module):
    def __init__(self, latent_size, img_size):
        super().__init__()
        l = latent_size
        self.main = nn.Sequential(*_mlp([l, 2 * l, 4 * l, 8 * l, img_size]), upper_softmax())

    def forward(self, input):
        return self.main(input)


class Encoder(nn.Module):
    def __init__(self, latent_size, img_size):
        super().__init__()
        l = latent_size
        self.main = _mlp([img_size, 8 * l, 4 * l, 2 * l, l])

    def forward(self, input):
        return self.main(input)


class Decoder(nn.Module):
    def __init__(self, latent_size, img_size):
        super().__init__()
        l = latent_size
        self.main = _mlp([l, 2 * l, 4 * l, 8 * l, img_size])

    def forward(self, input):
        return self.main(input)


class Detector(nn.Module):
    def __init__(self, latent_size, img_size, encoder, decoder):
        super().__init__()
        self.encoder = encoder(latent_size, img_size)
        self.decoder = decoder(latent_size, img_size)

    def forward(self, input):
        enc = self.encoder(input)
        dec = self.decoder(enc)
        return enc.view(input.size(0), -1), dec.view(input.size(0), -1)
