"""Drop-in ``RBF`` / ``MMDLossConstrained`` backed by the sm_100a kernels (C ABI).

Mirrors /root/reference/src/models/Mmd_loss_constrained.py: same class names, constructor
signatures, attributes (``bandwidth``, ``bandwidth_multipliers``, ``kernel``, ``weight``,
``device``) and semantics, including the ones that look like accidents:

* the bandwidth is computed on the first call and then frozen for the life of the ``RBF``
  object (``:16-22``), and ``kernel=RBF()`` is a default argument, i.e. ONE shared object
  for every ``MMDLossConstrained`` built without an explicit kernel (``:35``);
* ``s_m = bandwidth * multiplier_m`` is rounded to fp32 (``:26``);
* biased estimator with the diagonal included (``:46-49``);
* ``torch.topk(U, 1, 0)`` penalty with the lowest row index winning ties (``:50``).

One evaluation = ONE pass of the fused kernel: forward value and dL/dX, dL/dY are produced
together (the loss is linear in the kernel sums, so ``backward`` only scales the stashed
gradients by ``grad_output``).  There is no CPU implementation here.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch
from torch import nn

from . import _lib


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _as_input(t: torch.Tensor, device) -> torch.Tensor:
    # the reference casts the batch to float32 on cuda (src/vgan.py:602)
    return t.detach().to(device=device, dtype=torch.float32).contiguous()


def _is_ladder(mults) -> bool:
    m = [float(v) for v in mults]
    return len(m) == 5 and all(m[i + 1] == 2.0 * m[i] for i in range(4))


class _Workspace:
    """Device scratch for one problem shape, reused across calls."""

    def __init__(self, Nx, Ny, D, D_u, flags, device):
        self.layout = _lib.layout(Nx, Ny, D, D_u, flags)
        self.buf = torch.empty(self.layout.total + 1024, dtype=torch.uint8, device=device)
        base = self.buf.data_ptr()
        self.ptr = (base + 1023) // 1024 * 1024
        self.sums = torch.zeros(4, dtype=torch.float64, device=device)
        self.penalty = torch.zeros(1, dtype=torch.float32, device=device)
        self.loss = torch.zeros(1, dtype=torch.float32, device=device)
        self.inv_scales = torch.zeros(_lib.MAX_SCALES, dtype=torch.float32, device=device)


_ws_cache = {}


def _workspace(Nx, Ny, D, D_u, flags, device) -> _Workspace:
    key = (Nx, Ny, D, D_u, flags & (_lib.PREC_FP32 | _lib.BF16X3), str(device),
           torch.cuda.current_stream().cuda_stream)
    ws = _ws_cache.get(key)
    if ws is None:
        if len(_ws_cache) > 16:
            _ws_cache.clear()
        ws = _ws_cache[key] = _Workspace(Nx, Ny, D, D_u, flags, device)
    return ws


def default_precision(D: int) -> str:
    """bf16 tensor-core path by default; tiny feature widths (the VGAN encodings,
    latent = D//16) go to the fp32 CUDA-core path where a 64-wide K chunk would be padding."""
    return "fp32" if D < 32 else "bf16"


def _precision_flags(precision: str, D: int) -> int:
    """"bf16": bf16 operands on the tensor cores (2e-3).  "fp32": fp32-class results (1e-4) — split bf16 operands on
    the tensor cores ("bf16x3": 2.5-35x faster than the CUDA-core kernel from D = 16 to 4096, profiles/r1d_sweep.jsonl),
    the CUDA-core kernel ("simt") only for the tiny widths of the VGAN encodings (D < 32), where a 64-wide K chunk is
    mostly padding.  "simt" / "bf16x3" force one of the two."""
    if precision == "bf16":
        return _lib.PREC_BF16
    if precision == "simt" or (precision == "fp32" and D < 32):
        return _lib.PREC_FP32
    if precision in ("bf16x3", "fp32"):
        return _lib.PREC_BF16 | _lib.BF16X3
    raise RuntimeError(f"vgan_b200: unknown precision {precision!r} (bf16, fp32, bf16x3, simt)")


def mmd2_eval(X: torch.Tensor, Y: Optional[torch.Tensor], U: Optional[torch.Tensor], *,
              weight: float, bw_dev: torch.Tensor, frozen: bool, mults, precision: str = "bf16",
              need_dX: bool = False, need_dY: bool = True, masked: bool = False,
              penalty_grad: bool = True, row_begin: int = 0, row_end: Optional[int] = None,
              xx_part: int = 0, xx_nparts: int = 1, center: bool = True, finalize: bool = True,
              k1_events=None):
    """Run prologue -> scales -> fused fwd+bwd -> penalty -> finalize on the current stream.

    ``masked=True``: Y is formed as U*X inside the prologue and the returned ``dY`` is already
    dL/dU = dL/dY * X (+ penalty gradient).  Returns dict(loss, sums, dX, dY, penalty); tensors
    stay on the device, nothing synchronises."""
    L = _lib.lib()
    dev = X.device
    if dev.type != "cuda":
        raise RuntimeError("vgan_b200: inputs must live on a CUDA (sm_100) device; there is no CPU path")
    Nx, D = X.shape
    Ny = Nx if masked else Y.shape[0]
    D_u = 0 if U is None else U.shape[1]
    flags = _precision_flags(precision, D)
    if center:
        flags |= _lib.CENTER
    if need_dX:
        flags |= _lib.NEED_DX
    if not (flags & _lib.PREC_FP32) and _is_ladder(mults):
        flags |= _lib.LADDER
    if os.environ.get("VGAN_B200_NO_CLUSTER"):
        flags |= _lib.NO_CLUSTER
    if os.environ.get("VGAN_B200_NO_PAIR"):
        flags |= _lib.NO_PAIR
    ws = _workspace(Nx, Ny, D, D_u, flags, dev)
    st = _stream()
    row_end = max(Nx, Ny) if row_end is None else row_end
    n_rows = row_end - row_begin

    _lib.check(L.vgan_b200_prologue(_ptr(X), _ptr(U) if masked else None, None if masked else _ptr(Y),
                                    Nx, Ny, D, flags, ws.ptr, st), "prologue")
    m_arr = (C.c_float * len(mults))(*[float(v) for v in mults])
    _lib.check(L.vgan_b200_scales(ws.ptr, Nx, Ny, D, _ptr(bw_dev), 1 if frozen else 0, m_arr, len(mults),
                                  _ptr(ws.inv_scales), st), "scales")
    ws.sums.zero_()
    ws.penalty.zero_()
    want_grad = need_dY or need_dX
    dY = torch.empty((min(row_end, Ny) - row_begin, D), dtype=torch.float32, device=dev) if want_grad else None
    dX = torch.empty((min(row_end, Nx) - row_begin, D), dtype=torch.float32, device=dev) if need_dX else None
    mul = X[row_begin:] if (masked and want_grad) else None
    if k1_events is not None:
        k1_events[0].record()
    _lib.check(L.vgan_b200_mmd2_fwd_bwd(ws.ptr, Nx, Ny, D, row_begin, row_end, xx_part, xx_nparts,
                                        _ptr(ws.inv_scales), len(mults), flags, _ptr(ws.sums), _ptr(dY),
                                        _ptr(mul), _ptr(dX), st), "mmd2_fwd_bwd")
    if k1_events is not None:
        k1_events[1].record()
    dU_pen = None
    if U is not None and weight != 0:
        tgt = None
        if penalty_grad:
            if masked and dY is not None:
                tgt = dY                       # accumulate straight into dL/dU
            else:
                dU_pen = torch.zeros((n_rows if masked else U.shape[0], D_u), dtype=torch.float32, device=dev)
                tgt = dU_pen
        rb = row_begin if masked else 0
        nr = (tgt.shape[0] if tgt is not None else 0)
        _lib.check(L.vgan_b200_colmax_penalty(_ptr(U), U.shape[0], D_u, float(weight), ws.ptr + ws.layout.off_cmax,
                                              _ptr(ws.penalty), _ptr(tgt), rb, nr, st), "colmax_penalty")
    if finalize:
        _lib.check(L.vgan_b200_finalize_loss(_ptr(ws.sums), Nx, Ny, _ptr(ws.penalty), _ptr(ws.loss), st),
                   "finalize_loss")
    return dict(loss=ws.loss, sums=ws.sums, dX=dX, dY=dY, dU_penalty=dU_pen, penalty=ws.penalty, ws=ws)


class RBF(nn.Module):
    """Multi-bandwidth Gaussian kernel (Mmd_loss_constrained.py:5-26)."""

    def __init__(self, n_kernels=5, mul_factor=2.0, bandwidth=None):
        super().__init__()
        device = torch.device('cuda:0' if torch.cuda.is_available() else 'cpu')
        self.bandwidth_multipliers = (mul_factor ** (torch.arange(n_kernels) - n_kernels // 2)).to(device)
        self._mults = [float(mul_factor) ** (k - n_kernels // 2) for k in range(n_kernels)]
        self.bandwidth = bandwidth
        self._bw_dev = None            # device copy, the one the kernels read

    # -- frozen-bandwidth bookkeeping shared by RBF.forward and MMDLossConstrained.forward
    def _bw_state(self, device):
        frozen = self.bandwidth is not None
        if self._bw_dev is None or self._bw_dev.device != device:
            self._bw_dev = torch.zeros(1, dtype=torch.float32, device=device)
            if frozen:
                self._bw_dev.fill_(float(self.bandwidth))
        return self._bw_dev, frozen

    def _freeze(self):
        if self.bandwidth is None:
            self.bandwidth = self._bw_dev[0]       # 0-dim device tensor; .item() works, no sync here

    def reset_bandwidth(self):
        """Not in the reference: forget the frozen bandwidth (next call recomputes it)."""
        self.bandwidth = None
        self._bw_dev = None

    def get_bandwidth(self, L2_distances):
        if self.bandwidth is None:
            n_samples = L2_distances.shape[0]
            self.bandwidth = L2_distances.data.sum() / (n_samples ** 2 - n_samples)
            self._bw_dev = None
            return self.bandwidth
        return self.bandwidth

    def forward(self, X):
        """[n, n] kernel matrix, materialised (small n only — parity / GoF-test use)."""
        L = _lib.lib()
        dev = X.device
        if dev.type != "cuda":
            raise RuntimeError("vgan_b200.RBF: input must be on a CUDA (sm_100) device")
        Z = _as_input(X, dev)
        n, D = Z.shape
        bw, frozen = self._bw_state(dev)
        st = _stream()
        if not frozen:
            # closed form of sum_ij |zi-zj|^2 / (n^2-n)  (Mmd_loss_constrained.py:16-22), on device
            zs = Z.double()
            tot = 2.0 * n * (zs * zs).sum() - 2.0 * (zs.sum(0) ** 2).sum()
            bw.copy_((tot / (n * n - n)).float().reshape(1))
        inv = 1.0 / (bw * torch.tensor(self._mults, dtype=torch.float32, device=dev))
        K = torch.empty((n, n), dtype=torch.float32, device=dev)
        _lib.check(L.vgan_b200_mmd2_matrix(_ptr(Z), n, D, _ptr(inv), len(self._mults), 0.0, _ptr(K), st),
                   "mmd2_matrix")
        self._freeze()
        return K


class _MMD2Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, X, Y, U, module):
        dev = X.device
        Xc, Yc = _as_input(X, dev), _as_input(Y, dev)
        Uc = _as_input(U, dev) if (U is not None and module.weight != 0) else None
        kernel = module.kernel
        bw, frozen = kernel._bw_state(dev)
        prec = module.precision or default_precision(Xc.shape[1])
        need_dX = bool(ctx.needs_input_grad[0])
        need_dY = bool(ctx.needs_input_grad[1])
        need_dU = bool(ctx.needs_input_grad[2]) and Uc is not None
        out = mmd2_eval(Xc, Yc, Uc, weight=float(module.weight), bw_dev=bw, frozen=frozen,
                        mults=kernel._mults, precision=prec, need_dX=need_dX, need_dY=need_dY or need_dX,
                        masked=False, penalty_grad=need_dU, center=module.center)
        kernel._freeze()
        ctx.grads = (out["dX"] if need_dX else None, out["dY"] if need_dY else None,
                     out["dU_penalty"] if need_dU else None)
        ctx.in_dtypes = (X.dtype, Y.dtype, None if U is None else U.dtype)
        module.last_sums = out["sums"]
        return out["loss"][0].clone()

    @staticmethod
    def backward(ctx, g):
        dX, dY, dU = ctx.grads
        dts = ctx.in_dtypes
        return (None if dX is None else (dX * g).to(dts[0]),
                None if dY is None else (dY * g).to(dts[1]),
                None if dU is None else (dU * g).to(dts[2]), None)


class _MaskedFn(torch.autograd.Function):
    """loss(X, U*X, U) with the masking fused into the prologue and dL/dU written directly."""

    @staticmethod
    def forward(ctx, X, U, module):
        dev = X.device
        Xc, Uc = _as_input(X, dev), _as_input(U, dev)
        kernel = module.kernel
        bw, frozen = kernel._bw_state(dev)
        prec = module.precision or default_precision(Xc.shape[1])
        need_dU = bool(ctx.needs_input_grad[1])
        out = mmd2_eval(Xc, None, Uc, weight=float(module.weight), bw_dev=bw, frozen=frozen,
                        mults=kernel._mults, precision=prec, need_dX=False, need_dY=need_dU, masked=True,
                        penalty_grad=need_dU, center=module.center)
        kernel._freeze()
        ctx.dU = out["dY"] if need_dU else None
        ctx.u_dtype = U.dtype
        module.last_sums = out["sums"]
        return out["loss"][0].clone()

    @staticmethod
    def backward(ctx, g):
        return None, (None if ctx.dU is None else (ctx.dU * g).to(ctx.u_dtype)), None


class MMDLossConstrained(nn.Module):
    '''
    Constrained loss by the number of features selected — B200 kernel behind the reference
    interface (Mmd_loss_constrained.py:29-50).  Extra keyword-only knobs (defaults reproduce the
    reference): ``precision`` in {None (auto), "bf16", "fp32"}, ``center``.
    '''

    def __init__(self, weight, kernel=RBF(), *, precision=None, center=True):
        super().__init__()
        self.kernel = kernel
        self.weight = weight
        self.precision = precision
        self.center = center
        self.device = torch.device('cuda:0' if torch.cuda.is_available() else 'cpu')
        self.last_sums = None

    def forward(self, X, Y, U):
        if X.dim() != 2 or Y.dim() != 2 or X.shape[1] != Y.shape[1]:
            raise RuntimeError(f"MMDLossConstrained: X {tuple(X.shape)} and Y {tuple(Y.shape)} must be 2-D "
                               "with the same number of columns")
        if X.device.type != "cuda":
            raise RuntimeError("vgan_b200.MMDLossConstrained needs CUDA tensors on an sm_100 device "
                               "(no CPU fallback exists)")
        loss = _MMD2Fn.apply(X, Y, U, self)
        self.bandwidth = self.kernel.bandwidth
        self.bandwidth_multipliers = self.kernel.bandwidth_multipliers
        return loss

    def forward_masked(self, X, U):
        """Same value and gradient as ``forward(X, U*X, U)`` (src/vgan.py:615-616) in one fused call."""
        if X.shape != U.shape:
            raise RuntimeError("forward_masked: X and U must have the same shape")
        if X.device.type != "cuda":
            raise RuntimeError("vgan_b200.MMDLossConstrained needs CUDA tensors on an sm_100 device")
        loss = _MaskedFn.apply(X, U, self)
        self.bandwidth = self.kernel.bandwidth
        self.bandwidth_multipliers = self.kernel.bandwidth_multipliers
        return loss
