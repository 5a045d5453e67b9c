"""GPU: the bench shape itself (BASELINE.json configs[3]: N = 65536, D = 1024).  The full fp64 oracle would take
hours there, so parity is asserted (a) against the oracle's blockwise port on a SAMPLE of gradient rows versus all
2N columns, and (b) through size-independent properties: row shards add up to the unsharded result, an all-ones
mask makes the two samples identical (S_xx = S_xy = S_yy, zero gradient), and permuting the samples permutes the
gradient rows and leaves the sums alone."""
import numpy as np
import pytest
import torch

from oracle import mmd_oracle as orc

pytestmark = pytest.mark.gpu

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


def test_row_sample_matches_oracle_port(data):
    from vgan_b200 import RBF
    X, U = data
    Xd, Ud = X.cuda(), U.cuda()
    kernel = RBF()
    sums, dU = _eval(Xd, Ud, kernel)
    bw = float(kernel.bandwidth)
    r0, r1 = 33024, 33280                                       # two row tiles in the middle of the batch
    want_sums, want_dU = orc.cpu_blockwise_fwd_bwd(X, U, bw, r0, r1)
    # the same rows alone: S_xy, S_yy of these rows must match the port, and the rows of the full run
    s_rows, dU_rows = _eval(Xd, Ud, kernel, row_begin=r0, row_end=r1, xx_part=10 ** 6 - 1, xx_nparts=10 ** 6)
    assert abs(float(s_rows[1]) - want_sums[1]) <= 2e-3 * want_sums[1]
    assert abs(float(s_rows[2]) - want_sums[2]) <= 2e-3 * want_sums[2]
    err = float((dU_rows.double().cpu() - want_dU.double()).norm() / want_dU.double().norm())
    assert err <= 2e-2, err
    assert float((dU[r0:r1] - dU_rows).norm() / dU_rows.norm()) <= 1e-5


def test_fp32_class_row_sample_matches_oracle_port(data):
    """The split-operand (bf16x3) path at the bench shape: three gradient passes accumulating into the same rows
    through the persistent stream-K kernel; 1e-4 against the port's fp32 blockwise evaluation of the same rows."""
    from vgan_b200 import RBF
    X, U = data
    Xd, Ud = X.cuda(), U.cuda()
    kernel = RBF()
    sums, dU = _eval(Xd, Ud, kernel, precision="bf16x3")
    bw = float(kernel.bandwidth)
    r0, r1 = 33024, 33280
    want_sums, want_dU = orc.cpu_blockwise_fwd_bwd(X, U, bw, r0, r1)
    s_rows, dU_rows = _eval(Xd, Ud, kernel, precision="bf16x3", row_begin=r0, row_end=r1, xx_part=10 ** 6 - 1,
                            xx_nparts=10 ** 6)
    assert abs(float(s_rows[1]) - want_sums[1]) <= 1e-4 * want_sums[1]
    assert abs(float(s_rows[2]) - want_sums[2]) <= 1e-4 * want_sums[2]
    err = float((dU_rows.double().cpu() - want_dU.double()).norm() / want_dU.double().norm())
    assert err <= 1e-4, err
    assert float((dU[r0:r1] - dU_rows).norm() / dU_rows.norm()) <= 1e-5
    # and the bf16 path agrees with it to bf16 accuracy
    sums16, dU16 = _eval(Xd, Ud, kernel)
    assert torch.allclose(sums16[:3], sums[:3], rtol=2e-3)
    assert float((dU16 - dU).norm() / dU.norm()) <= 2e-2


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
