"""GPU: the bench shape itself (BASELINE.json configs[3]: N = 65536, D = 1024).  The full fp64 oracle would take
hours there, so parity is asserted (a) against the oracle's blockwise port on a SAMPLE of gradient rows versus all
2N columns, and (b) through size-independent properties: row shards add up to the unsharded result, an all-ones
mask makes the two samples identical (S_xx = S_xy = S_yy, zero gradient), and permuting the samples permutes the
gradient rows and leaves the sums alone."""
import numpy as np
import pytest
import torch

import parity_util as pu
from oracle import mmd_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, scope="module")
def _write_parity_report():
    yield
    pu.flush()

N, D = 65536, 1024


@pytest.fixture(scope="module")
def data():
    g = torch.Generator().manual_seed(0)
    X = torch.randn(N, D, generator=g)
    p = torch.softmax(torch.randn(N, D, generator=g), 1)
    U = torch.where(p < 1.0 / D, p, torch.ones_like(p))          # upper_softmax (Generator.py:18-22)
    return X, U


def _eval(Xd, Ud, kernel, precision="bf16", **kw):
    from vgan_b200.loss import mmd2_eval
    bw, frozen = kernel._bw_state(Xd.device)
    out = mmd2_eval(Xd, None, Ud, weight=0.0, bw_dev=bw, frozen=frozen, mults=kernel._mults, precision=precision,
                    masked=True, **kw)
    kernel._freeze()
    torch.cuda.synchronize()
    return out["sums"].clone(), out["dY"].clone()


R0, R1 = 33024, 33280                                       # two row tiles in the middle of the batch


@pytest.fixture(scope="module")
def truth(data):
    """fp64 truth for the sampled gradient rows and the S_xx row tile (oracle.bench_sample_fp64), per bandwidth."""
    cache = {}

    def get(bw):
        key = round(bw, 3)
        if key not in cache:
            cache[key] = orc.bench_sample_fp64(data[0], data[1], bw, R0, R1, R0 // 128)
        return cache[key]
    return get


@pytest.mark.parametrize("prec,tol", [("bf16", 2e-3), ("bf16x3", 1e-4)])
def test_row_sample_matches_oracle(data, truth, prec, tol):
    """256 gradient rows against all 2N columns (S_xy, S_yy share and dL/dU rows: the pair kernel) AND the S_xx share of
    one X row tile walking the upper triangle (mmd2_xx_kernel) against the fp64 oracle: 2e-3 on the bf16-operand
    path, 1e-4 on the split-operand (fp32-class) path — sums and gradient alike."""
    from vgan_b200 import RBF
    X, U = data
    Xd, Ud = X.cuda(), U.cuda()
    kernel = RBF()
    sums, dU = _eval(Xd, Ud, kernel, precision=prec)
    want = truth(float(kernel.bandwidth))
    # the same rows alone: S_xy, S_yy of these rows and the S_xx tile must match the oracle, and the rows of the full run
    s_rows, dU_rows = _eval(Xd, Ud, kernel, precision=prec, row_begin=R0, row_end=R1, xx_part=R0 // 128, xx_nparts=10 ** 6)
    e = dict(sxy_rows=abs(float(s_rows[1]) - want["sxy"]) / want["sxy"],
             syy_rows=abs(float(s_rows[2]) - want["syy"]) / want["syy"],
             sxx_tile=abs(float(s_rows[0]) - want["sxx_tile"]) / want["sxx_tile"],
             dU=pu.rel_l2(dU_rows, want["dU"]), ref_fp32_dU=want["ref_fp32_dU"])
    pu.record(f"bench_shape/N{N}_D{D}/{prec}", N_x=N, N_y=N, D=D, rows=R1 - R0, **e)
    assert e["sxy_rows"] <= tol and e["syy_rows"] <= tol and e["sxx_tile"] <= tol, e
    assert e["dU"] <= tol, e
    assert float((dU[R0:R1] - dU_rows).norm() / dU_rows.norm()) <= 1e-5


def test_bf16_agrees_with_fp32_class(data):
    from vgan_b200 import RBF
    X, U = data
    Xd, Ud = X.cuda(), U.cuda()
    kernel = RBF()
    sums, dU = _eval(Xd, Ud, kernel, precision="bf16x3")
    sums16, dU16 = _eval(Xd, Ud, kernel)
    assert torch.allclose(sums16[:3], sums[:3], rtol=2e-3)
    assert float((dU16 - dU).norm() / dU.norm()) <= 2e-3


def test_row_shards_add_up(data):
    from vgan_b200 import RBF
    X, U = data
    Xd, Ud = X.cuda(), U.cuda()
    kernel = RBF()
    sums, dU = _eval(Xd, Ud, kernel)
    acc = torch.zeros(3, dtype=torch.float64, device="cuda")
    for world in (2, 8):
        acc.zero_()
        for rank in range(world):
            b, e = rank * N // world, (rank + 1) * N // world
            s, g = _eval(Xd, Ud, kernel, row_begin=b, row_end=e, xx_part=rank, xx_nparts=world)
            acc += s[:3]
            assert float((g - dU[b:e]).norm() / dU[b:e].norm()) <= 1e-5
        assert torch.allclose(acc, sums[:3], rtol=1e-6)


def test_identity_mask_gives_zero_mmd(data):
    from vgan_b200 import RBF
    X, _ = data
    Xd = X.cuda()
    ones = torch.ones_like(Xd)
    sums, dU = _eval(Xd, ones, RBF())
    sxx, sxy, syy = (float(v) for v in sums[:3])
    assert abs(sxy - sxx) <= 1e-6 * sxx and abs(syy - sxx) <= 1e-6 * sxx
    _, dU_rand = _eval(Xd, torch.rand_like(Xd), RBF())
    assert float(dU.norm()) <= 1e-3 * float(dU_rand.norm())


def test_permutation_equivariance(data):
    from vgan_b200 import RBF
    X, U = data
    Xd, Ud = X.cuda(), U.cuda()
    kernel = RBF()
    sums, dU = _eval(Xd, Ud, kernel)
    perm = torch.randperm(N, generator=torch.Generator().manual_seed(1)).cuda()
    sums_p, dU_p = _eval(Xd[perm].contiguous(), Ud[perm].contiguous(), kernel)
    assert torch.allclose(sums_p[:3], sums[:3], rtol=2e-6)
    assert float((dU_p - dU[perm]).norm() / dU.norm()) <= 1e-4
