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
from typing import Optional

import torch
from torch import nn

from . import _lib


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _ctx(device: torch.device):
    return _lib.ctx(device.index if device.index is not None else torch.cuda.current_device())


def _as_input(t: torch.Tensor, device) -> torch.Tensor:
    # the reference casts the batch to float32 on cuda (src/vgan.py:602)
    return t.detach().to(device=device, dtype=torch.float32).contiguous()


def _is_ladder(mults) -> bool:
    m = [float(v) for v in mults]
    return len(m) == 5 and all(m[i + 1] == 2.0 * m[i] for i in range(4))


class _Workspace:
    """Device scratch for one problem shape, reused across calls; typed views of the parts the host side touches
    (results, and on several GPUs the blocks that are exchanged — see dist.py)."""

    def __init__(self, Nx, Ny, D, D_u, mults, flags, device):
        self.layout = L = _lib.layout(Nx, Ny, D, D_u, len(mults), flags)
        self.buf = torch.empty(L.total + 1024, dtype=torch.uint8, device=device)
        base = self.buf.data_ptr()
        self.ptr = (base + 1023) // 1024 * 1024
        self._off0 = self.ptr - base
        self.mults_c = (C.c_float * len(mults))(*[float(v) for v in mults])
        self.sums = self.view(L.off_sums, 4, torch.float64)
        self.colsum = self.view(L.off_colsum, L.Dp, torch.float64)[:D]
        self.reduced = self.view(L.off_sums, 4 + L.Dp, torch.float64)         # [S_xx, S_xy, S_yy, spare | colsum]
        self.stats = self.view(L.off_stats, 2 + 2 * L.Dp, torch.float64)
        self.cmax = self.view(L.off_cmax, D_u, torch.int64) if D_u > 0 else None
        self.penalty = self.view(L.off_misc, 1, torch.float32)
        self.loss = self.view(L.off_misc + 4, 1, torch.float32)
        rows = L.Nxp + L.Nyp
        fp32 = (flags & _lib.PREC_FP32) != 0
        width = L.Dp * (3 if (flags & _lib.BF16X3) else 1)
        if fp32:
            self.operand = [self.view(L.off_zf, rows * L.Dp, torch.float32).view(rows, L.Dp)]
            self.norms = self.view(L.off_nrm, rows, torch.float32)
        else:
            n_arr = 2 if (flags & _lib.BF16X3) else 1
            z = self.view(L.off_zq, n_arr * rows * width, torch.bfloat16).view(n_arr, rows, width)
            self.operand = [z[k] for k in range(n_arr)]
            self.norms = self.view(L.off_nrms, rows, torch.float32)
        self.layout_ref = C.byref(self.layout)

    def view(self, off: int, count: int, dtype: torch.dtype) -> torch.Tensor:
        nbytes = count * torch.empty((), dtype=dtype).element_size()
        return self.buf[self._off0 + off:self._off0 + off + nbytes].view(dtype)


_ws_cache = {}


def _workspace(Nx, Ny, D, D_u, mults, flags, device) -> _Workspace:
    key = (Nx, Ny, D, D_u, tuple(float(v) for v in mults), flags, str(device),
           torch.cuda.current_stream(device).cuda_stream)
    ws = _ws_cache.get(key)
    if ws is None:
        if len(_ws_cache) > 16:
            _ws_cache.clear()
        ws = _ws_cache[key] = _Workspace(Nx, Ny, D, D_u, mults, flags, device)
    return ws


def default_precision(D: int) -> str:
    """bf16 tensor-core path by default (bf16 operands meet the stated 2e-3 from D = 32); narrower inputs (the VGAN
    encodings, latent = D//16) take the fp32-class path, which is the split-operand tensor-core kernel as well."""
    return "fp32" if D < 32 else "bf16"


def _precision_flags(precision: str, D: int) -> int:
    """"bf16": bf16 operands on the tensor cores (2e-3).  "fp32": fp32-class results (1e-4) — split bf16 operands on
    the tensor cores ("bf16x3") at every width (faster than the CUDA-core kernel wherever both were measured).
    "simt" forces the CUDA-core fp32 kernel, "bf16x3" the split-operand one."""
    if precision == "bf16":
        return _lib.PREC_BF16
    if precision == "simt":
        return _lib.PREC_FP32
    if precision in ("bf16x3", "fp32"):
        return _lib.PREC_BF16 | _lib.BF16X3
    raise RuntimeError(f"vgan_b200: unknown precision {precision!r} (bf16, fp32, bf16x3, simt)")


def make_flags(precision: str, D: int, mults, center: bool = True, need_dX: bool = False) -> int:
    flags = _precision_flags(precision, D)
    if center:
        flags |= _lib.CENTER
    if need_dX:
        flags |= _lib.NEED_DX
    if not (flags & _lib.PREC_FP32) and _is_ladder(mults):
        flags |= _lib.LADDER
    return flags


def mmd2_eval(X: torch.Tensor, Y: Optional[torch.Tensor], U: Optional[torch.Tensor], *,
              weight: float, bw_dev: torch.Tensor, frozen: bool, mults, precision: str = "bf16",
              need_dX: bool = False, need_dY: bool = True, masked: bool = False,
              penalty_grad: bool = True, row_begin: int = 0, row_end: Optional[int] = None,
              xx_part: int = 0, xx_nparts: int = 1, center: bool = True, k1_events=None):
    """stats -> scales -> pack -> fused fwd+bwd -> finalize on the current stream of X's device.

    ``masked=True``: Y is formed as U*X inside the prologue and the returned ``dY`` is already
    dL/dU = dL/dY * X (+ penalty gradient).  A row range / S_xx part selects a shard of K1 (the prologue still covers
    every row: this is the single-process form; dist.py splits the prologue as well).
    Returns dict(loss, sums, colsum, dX, dY, dU_penalty, penalty, ws); tensors stay on the device, nothing
    synchronises.  ``loss``, ``sums``, ``colsum`` and ``penalty`` are views of the workspace (valid until the next
    evaluation of the same shape on the same stream)."""
    L = _lib.lib()
    dev = X.device
    if dev.type != "cuda":
        raise RuntimeError("vgan_b200: inputs must live on a CUDA (sm_100) device; there is no CPU path")
    Nx, D = X.shape
    Ny = Nx if masked else Y.shape[0]
    use_pen = U is not None and weight != 0
    D_u = U.shape[1] if use_pen else 0
    flags = make_flags(precision, D, mults, center, need_dX)
    with torch.cuda.device(dev):
        ws = _workspace(Nx, Ny, D, D_u, mults, flags, dev)
        ctx, st = _ctx(dev), _stream(dev)
        n_all = max(Nx, Ny)
        row_end = n_all if row_end is None else row_end
        want_grad = need_dY or need_dX
        dY = torch.empty((max(min(row_end, Ny) - row_begin, 0), D), dtype=torch.float32, device=dev) if want_grad else None
        dX = torch.empty((max(min(row_end, Nx) - row_begin, 0), D), dtype=torch.float32, device=dev) if need_dX else None
        if dY is not None and dY.shape[0] == 0 and not need_dX:
            dY = None                          # empty shard: forward sums of its S_xx tiles only
        mul = X[row_begin:] if (masked and dY is not None) else None
        dU_pen, tgt = None, None
        if use_pen and penalty_grad:
            if masked and dY is not None:
                tgt = dY                       # the penalty gradient goes straight into dL/dU
            else:
                dU_pen = torch.zeros((row_end - row_begin if masked else U.shape[0], D_u), dtype=torch.float32, device=dev)
                tgt = dU_pen
        Up = _ptr(U) if (masked or use_pen) else None
        Yp = None if masked else _ptr(Y)
        whole = row_begin == 0 and row_end >= n_all and xx_nparts == 1 and k1_events is None
        if whole:
            args = _lib.EvalArgs(C.pointer(ws.layout), ws.ptr, _ptr(X), Up, Yp, _ptr(bw_dev), 1 if frozen else 0,
                                 ws.mults_c, float(weight), _ptr(dY), _ptr(mul), _ptr(dX), _ptr(tgt), None, st)
            _lib.check(L.vgan_b200_mmd2_eval(ctx, C.byref(args)), "mmd2_eval")
        else:
            lay = ws.layout_ref
            _lib.check(L.vgan_b200_stats(ctx, lay, ws.ptr, _ptr(X), Up, Yp, 0, n_all, st), "stats")
            _lib.check(L.vgan_b200_scales(ctx, lay, ws.ptr, _ptr(bw_dev), 1 if frozen else 0, ws.mults_c, float(weight),
                                          st), "scales")
            _lib.check(L.vgan_b200_pack(ctx, lay, ws.ptr, _ptr(X), Up, Yp, 0, n_all, st), "pack")
            if k1_events is not None:
                k1_events[0].record()
            _lib.check(L.vgan_b200_mmd2_fwd_bwd(ctx, lay, ws.ptr, _ptr(X), Up, Yp, row_begin, row_end, xx_part, xx_nparts,
                                                _ptr(dY), _ptr(mul), _ptr(dX), st), "mmd2_fwd_bwd")
            if k1_events is not None:
                k1_events[1].record()
            rb = row_begin if masked else 0
            nr = tgt.shape[0] if tgt is not None else 0
            _lib.check(L.vgan_b200_finalize(ctx, lay, ws.ptr, float(weight), _ptr(tgt), rb, nr,
                                            1 if (tgt is not None and tgt is dY) else 0, None, st), "finalize")
    return dict(loss=ws.loss, sums=ws.sums, colsum=ws.colsum, dX=dX, dY=dY, dU_penalty=dU_pen, penalty=ws.penalty, ws=ws)


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
        with torch.cuda.device(dev):
            Z = _as_input(X, dev)
            n, D = Z.shape
            bw, frozen = self._bw_state(dev)
            st = _stream(dev)
            if not frozen:
                # closed form of sum_ij |zi-zj|^2 / (n^2-n)  (Mmd_loss_constrained.py:16-22), on device
                zs = Z.double()
                tot = 2.0 * n * (zs * zs).sum() - 2.0 * (zs.sum(0) ** 2).sum()
                bw.copy_((tot / (n * n - n)).float().reshape(1))
            inv = 1.0 / (bw * torch.tensor(self._mults, dtype=torch.float32, device=dev))
            K = torch.empty((n, n), dtype=torch.float32, device=dev)
            _lib.check(L.vgan_b200_mmd2_matrix(_ctx(dev), _ptr(Z), n, D, _ptr(inv), len(self._mults), 0.0, _ptr(K), st),
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
        if ctx.needs_input_grad[0]:
            raise RuntimeError("forward_masked: a gradient with respect to X is not produced on the fused path; "
                               "call forward(X, U*X, U) when X requires grad")
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
        module.last_colsum = out["colsum"]
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
        self.last_colsum = None

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
        """Same value and dL/dU as ``forward(X, U*X, U)`` (src/vgan.py:615-616) in one fused call.  X must not
        require grad (the reference's no-KL step differentiates with respect to U only)."""
        if X.shape != U.shape:
            raise RuntimeError("forward_masked: X and U must have the same shape")
        if X.device.type != "cuda":
            raise RuntimeError("vgan_b200.MMDLossConstrained needs CUDA tensors on an sm_100 device")
        loss = _MaskedFn.apply(X, U, self)
        self.bandwidth = self.kernel.bandwidth
        self.bandwidth_multipliers = self.kernel.bandwidth_multipliers
        return loss
