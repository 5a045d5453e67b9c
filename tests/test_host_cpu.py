"""CPU-side checks: the C-ABI library builds, loads and exports every symbol declared in
include/vgan_b200.h; host classes keep the reference's surface; no compute without a GPU."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_build_and_exports_match_header():
    import __graft_entry__ as ge
    ge.build()
    from vgan_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    header = open(os.path.join(ROOT, "include", "vgan_b200.h")).read()
    declared = set(re.findall(r"\b(vgan_b200_[a-z0-9_]+)\s*\(", header))
    assert declared == set(_lib.EXPORTS)
    for name in declared:
        assert getattr(lib, name) is not None
    # the product library exports exactly the documented entry points: no probes, no debug switches, no C++ symbols
    import subprocess
    nm = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = {ln.split()[-1] for ln in nm.splitlines() if " T " in ln}
    assert exported == declared, exported ^ declared
    # the measurement entry points live in the debug build only
    dbg = ctypes.CDLL(_lib.DBG_LIB_PATH)
    for name in _lib.DEBUG_EXPORTS:
        assert getattr(dbg, name) is not None and name not in exported


def test_context_fails_loudly_without_device():
    """No GPU here: creating a context must fail with ERR_DEVICE and a message, never fall back."""
    from vgan_b200 import _lib
    if torch.cuda.is_available():
        pytest.skip("needs a machine without CUDA devices")
    h = ctypes.c_void_p()
    assert _lib.lib().vgan_b200_create(0, ctypes.byref(h)) == -2 and not h
    assert b"no CUDA device" in _lib.lib().vgan_b200_last_error()
    with pytest.raises(RuntimeError, match="no CUDA device"):
        _lib.ctx(0)


def test_layout_is_pure_host_logic():
    from vgan_b200 import _lib
    L = _lib.layout(500, 500, 20, 20, 5, _lib.PREC_BF16)
    assert (L.Dp, L.Nxp, L.Nyp) == (64, 512, 512)
    assert L.off_zq % 1024 == 0 and L.total >= L.off_zq + 1024 * 64 * 2
    L32 = _lib.layout(500, 500, 20, 20, 5, _lib.PREC_FP32)
    assert L32.off_nrm == L.off_nrm and L32.off_stats == L.off_stats and L32.off_zf == L.off_zq
    # one all-reduce serves [S_xx, S_xy, S_yy, - | colsum]: the column sums sit directly behind the four sums, and
    # everything vgan_b200_stats zeroes is one contiguous block
    assert L.off_colsum == L.off_sums + 32 and L.off_acc == L.off_sums
    assert L.off_acc + L.acc_bytes == L.off_misc + 64 and L.off_stats > L.off_colsum and L.off_cmax > L.off_stats
    # VGAN_B200_ONES (output flag): three constant-1 features where the padding has room and mmd_tc.cu keeps the row
    # operand resident — plain bf16 up to D_p = 256, split operands at D_p = 64; never on input, never for fp32 / D > 256
    ONES = 65536
    assert L.flags & ONES and not (L32.flags & ONES)
    assert _lib.layout(64, 64, 61, 0, 5, _lib.PREC_BF16).flags & ONES
    L62 = _lib.layout(64, 64, 62, 0, 5, _lib.PREC_BF16)                               # 62 + 3 > 64: one more chunk
    assert L62.Dp == 128 and L62.flags & ONES
    L62s = _lib.layout(64, 64, 62, 0, 5, _lib.PREC_BF16 | _lib.BF16X3)                # split operands: no room, no ones
    assert L62s.Dp == 64 and not (L62s.flags & ONES)
    assert not (_lib.layout(64, 64, 128, 0, 5, _lib.PREC_BF16 | ONES).flags & ONES)   # no padding; the input bit is ignored
    assert _lib.layout(64, 64, 20, 0, 5, _lib.PREC_BF16 | _lib.BF16X3).flags & ONES
    assert not (_lib.layout(64, 64, 100, 0, 5, _lib.PREC_BF16 | _lib.BF16X3).flags & ONES)   # split, D_p = 128: streamed
    assert _lib.layout(64, 64, 250, 0, 5, _lib.PREC_BF16).flags & ONES
    assert not (_lib.layout(64, 64, 300, 0, 5, _lib.PREC_BF16).flags & ONES)
    wide = _lib.layout(4096, 4096, 784, 784, 5, _lib.PREC_BF16 | _lib.LADDER)
    assert wide.Dp == 896 and wide.off_zq - wide.off_xchg == 256 * 128 * 1024
    with pytest.raises(RuntimeError):
        _lib.layout(0, 1, 1, 0, 5, 0)
    with pytest.raises(RuntimeError):
        _lib.layout(8, 8, 4, 0, 3, _lib.LADDER)          # the squaring ladder needs exactly 5 scales


def test_no_cpu_fallback():
    from vgan_b200 import MMDLossConstrained, RBF, MMDStatistic
    with pytest.raises(RuntimeError, match="CUDA"):
        MMDLossConstrained(0, kernel=RBF())(torch.randn(8, 4), torch.randn(8, 4), torch.rand(8, 4))
    with pytest.raises(RuntimeError, match="CUDA"):
        RBF()(torch.randn(8, 4))
    with pytest.raises(RuntimeError, match="CUDA"):
        MMDStatistic(4, 4)(torch.randn(4, 2), torch.randn(4, 2), alphas=[1.0])


def test_reference_surface():
    """Same constructor signatures / attributes as Mmd_loss_constrained.py:7,35."""
    import inspect
    from vgan_b200 import MMDLossConstrained, RBF
    sig = inspect.signature(RBF.__init__)
    assert list(sig.parameters)[1:] == ["n_kernels", "mul_factor", "bandwidth"]
    assert sig.parameters["n_kernels"].default == 5 and sig.parameters["mul_factor"].default == 2.0
    k = RBF()
    assert k.bandwidth is None and k.bandwidth_multipliers.tolist() == [0.25, 0.5, 1.0, 2.0, 4.0]
    sig = inspect.signature(MMDLossConstrained.__init__)
    assert list(sig.parameters)[1:3] == ["weight", "kernel"]
    a, b = MMDLossConstrained(0), MMDLossConstrained(10)
    assert a.kernel is b.kernel            # the shared default RBF() of the reference
    assert hasattr(a, "device") and a.weight == 0


def test_shard_rows():
    from vgan_b200.dist import shard_rows
    N = 65536
    parts = [shard_rows(N, r, 8) for r in range(8)]
    assert parts[0] == (0, 8192) and parts[-1] == (57344, 65536)
    assert all(b % 128 == 0 for b, _ in parts)
    covered = sum(e - b for b, e in parts)
    assert covered == N
    parts = [shard_rows(1000, r, 3) for r in range(3)]
    assert parts == [(0, 384), (384, 768), (768, 1000)]


def test_product_code_does_not_import_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "vgan_b200")):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f


def test_torch_ops_registered_cuda_only():
    """torch.ops.vgan_b200.* exist with their schemas, and the dispatcher has no CPU kernel for them."""
    import vgan_b200
    from vgan_b200 import ops
    for name in ops.OPS:
        assert hasattr(torch.ops.vgan_b200, name)
    bw = torch.zeros(1)
    with pytest.raises((NotImplementedError, RuntimeError)):
        torch.ops.vgan_b200.mmd2_masked_fwd_bwd(torch.randn(8, 4), torch.rand(8, 4), bw, False,
                                                [0.25, 0.5, 1.0, 2.0, 4.0], 10.0, "bf16", True)
    with pytest.raises((NotImplementedError, RuntimeError)):
        torch.ops.vgan_b200.mmd2_matrix(torch.randn(8, 4), torch.ones(1), 0.0)


def test_bench_reference_arm_line():
    """bench.py --impl reference (the CPU port of the reference path) prints the contract's JSON line; under torchrun's
    OMP_NUM_THREADS=1 it still takes every host thread."""
    import json
    import subprocess
    import sys
    env = dict(os.environ, OMP_NUM_THREADS="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--n", "2048",
                          "--d", "64", "--steps", "1", "--warmup", "0", "--ref-rows", "512"],
                         capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "evals/s" and line["higher_is_better"] is True
    assert line["metric"] == "mmd2_fwd_bwd_evals_per_sec" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["value"] == line["value"]
    assert line["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))
    assert line["e2e"] == {"value": line["value"], "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
