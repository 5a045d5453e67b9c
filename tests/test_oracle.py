"""CPU: the oracle restatement vs the golden vectors produced by the reference module
(oracle/make_fixtures.py) and, when /root/reference is mounted, vs the live module."""
import os

import numpy as np
import pytest
import torch

from oracle import mmd_oracle as orc
from oracle.make_fixtures import CASES, GOLDEN, build_case

RTOL = 1e-9


def _golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


@pytest.mark.parametrize("name", list(CASES))
def test_oracle_matches_golden(name):
    g = _golden(name)
    X, Y, U, weight, kind = build_case(name)
    np.testing.assert_allclose([X.sum(), Y.sum(), U.sum()], g["in_checksum"], rtol=1e-12)
    out = orc.loss_and_grads(X, Y, U, weight, bandwidth=None, need_dX=True, block=97)
    assert out["bandwidth"] == pytest.approx(float(g["bandwidth"]), rel=1e-11)
    assert orc.bandwidth_closed_form(np.vstack([X, Y])) == pytest.approx(float(g["bandwidth"]), rel=1e-10)
    assert out["loss"] == pytest.approx(float(g["loss"]), rel=RTOL, abs=1e-12)
    for key, val in (("sxx", out["sxx"]), ("sxy", out["sxy"]), ("syy", out["syy"])):
        assert val == pytest.approx(float(g[key]), rel=1e-10)
    if kind == "free":
        grads = {"dX": out["dX"].numpy(), "dY": out["dY"].numpy(), "dU": out["dU_penalty"].numpy()}
    else:
        grads = {"dX": out["dX"].numpy(), "dU": (out["dY"] * torch.tensor(X) + out["dU_penalty"]).numpy()}
    for key, val in grads.items():
        if key in g.files:
            np.testing.assert_allclose(val, g[key], rtol=1e-7, atol=1e-13)
        else:
            np.testing.assert_allclose(val[:24], g[key + "_head"], rtol=1e-7, atol=1e-13)
            np.testing.assert_allclose(val.sum(0), g[key + "_colsum"], rtol=1e-6, atol=1e-12)
            assert np.sqrt((val * val).sum()) == pytest.approx(float(g[key + "_l2"]), rel=1e-8)


@pytest.mark.parametrize("name", ["ragged_37x5", "cfg1_500x20"])
def test_frozen_bandwidth_second_call(name):
    """Mmd_loss_constrained.py:16-22: the bandwidth of call 1 is reused for call 2."""
    g = _golden(name)
    X, Y, U, weight, kind = build_case(name)
    assert float(g["bandwidth_after_second_call"]) == float(g["bandwidth"])
    out = orc.loss_and_grads(X * 1.5, Y * 1.5, U, weight, bandwidth=float(g["bandwidth"]), need_dX=False)
    assert out["loss"] == pytest.approx(float(g["loss_frozen_bw_scaled"]), rel=RTOL)


def test_rbf_matrix_and_multipliers():
    np.testing.assert_array_equal(orc.bandwidth_multipliers(), [0.25, 0.5, 1.0, 2.0, 4.0])
    rs = np.random.RandomState(3)
    Z = rs.randn(50, 6)
    bw = orc.bandwidth_direct(Z)
    K = orc.rbf_matrix(Z, bw)
    assert torch.allclose(torch.diag(K), torch.full((50,), 5.0, dtype=torch.float64))
    sxx, sxy, syy, _, _ = orc.mmd2_terms(Z[:20], Z[20:], bw, need_dX=False)
    assert sxx == pytest.approx(float(K[:20, :20].sum()), rel=1e-12)
    assert sxy == pytest.approx(float(K[:20, 20:].sum()), rel=1e-12)
    assert syy == pytest.approx(float(K[20:, 20:].sum()), rel=1e-12)


def test_ladder_identity():
    """exp(-d/(s*2^m)) for m=-2..2 is the squaring ladder t, t^2, ... of t = exp(-d/(4s))."""
    d = np.linspace(0, 40, 101)
    s = 3.7
    t = np.exp(-d / (4 * s))
    direct = sum(np.exp(-d / (s * m)) for m in orc.bandwidth_multipliers())
    ladder = t + t ** 2 + t ** 4 + t ** 8 + t ** 16
    np.testing.assert_allclose(ladder, direct, rtol=1e-13)


def test_colmax_penalty_ties_lowest_row():
    U = np.array([[0.2, 1.0, 0.5], [1.0, 1.0, 0.5], [1.0, 0.3, 0.1]])
    val, g = orc.colmax_penalty(U, 10.0)
    assert val == pytest.approx(10.0 * (0 + 0 + 0.5) / 3)
    expect = np.zeros((3, 3))
    expect[1, 0] = expect[0, 1] = expect[0, 2] = -10.0 / 3
    np.testing.assert_allclose(g.numpy(), expect)


def test_mmd_statistic_and_pval():
    rs = np.random.RandomState(0)
    a, b = rs.randn(40, 3), rs.randn(50, 3) + 0.8
    mmd, K = orc.mmd_statistic(a, b, [0.5])
    # direct double loop
    Z = np.vstack([a, b])
    d2 = ((Z[:, None] - Z[None]) ** 2).sum(-1) + 1e-5
    Kd = np.exp(-0.5 * d2)
    np.testing.assert_allclose(K.numpy(), Kd, rtol=1e-9)
    k1, k2, k12 = Kd[:40, :40], Kd[40:, 40:], Kd[:40, 40:]
    ref = (k1.sum() - np.trace(k1)) / (40 * 39) + (k2.sum() - np.trace(k2)) / (50 * 49) - 2 * k12.sum() / 2000
    assert mmd == pytest.approx(ref, rel=1e-10)
    perms = np.stack([rs.permutation(90) for _ in range(200)])
    p = orc.permutation_pval(K, 40, 50, perms)
    assert 0.0 <= p <= 0.05          # clearly different distributions
    p_same = orc.permutation_pval(orc.mmd_statistic(a, rs.randn(50, 3), [0.5])[1], 40, 50, perms)
    assert p_same > 0.05


def test_cpu_blockwise_baseline_matches_oracle():
    rs = np.random.RandomState(1)
    X = torch.tensor(rs.randn(300, 16), dtype=torch.float32)
    U = torch.tensor(rs.rand(300, 16), dtype=torch.float32)
    bw = orc.bandwidth_direct(torch.cat([X, U * X]))
    ref = orc.masked_loss_and_grad(X, U, 0.0, bandwidth=bw)
    sums, dU = orc.cpu_blockwise_fwd_bwd(X, U, bw, 0, 300, block=128)
    assert sums[0] == pytest.approx(ref["sxx"], rel=1e-5)
    assert sums[1] == pytest.approx(ref["sxy"], rel=1e-5)
    assert sums[2] == pytest.approx(ref["syy"], rel=1e-5)
    np.testing.assert_allclose(dU.numpy(), ref["dU"].numpy(), rtol=2e-3, atol=1e-8)


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/models"), reason="reference not mounted")
def test_oracle_vs_live_reference_fp64():
    from oracle.make_fixtures import run_reference
    out = run_reference("unequal_200_328")
    g = _golden("unequal_200_328")
    assert out["loss"] == pytest.approx(float(g["loss"]), rel=1e-13)
    np.testing.assert_allclose(out["dY"], g["dY"], rtol=1e-12)


@pytest.mark.parametrize("name", ["tiny_8x4", "ragged_37x5", "cfg1_500x20", "unequal_200_328", "dups_96x12"])
def test_torch_eager_matches_golden(name):
    """oracle/torch_eager.py (the torch-eager arm of tools/sweep.py) reproduces the reference module's fp64 results."""
    from oracle import torch_eager
    g = _golden(name)
    X, Y, U, weight, kind = build_case(name)
    Xt = torch.tensor(X, dtype=torch.float64, requires_grad=True)
    Ut = torch.tensor(U, dtype=torch.float64, requires_grad=True)
    Yt = torch.tensor(Y, dtype=torch.float64) if kind == "free" else Ut * Xt.detach()
    fn = torch_eager.MMDLossConstrained(weight, kernel=torch_eager.RBF(device="cpu"), device="cpu")
    loss = fn(Xt, Yt, Ut)
    loss.backward()
    assert float(fn.bandwidth) == pytest.approx(float(g["bandwidth"]), rel=1e-10)
    assert float(loss) == pytest.approx(float(g["loss"]), rel=1e-8, abs=1e-12)
    if "dU" in g.files:
        np.testing.assert_allclose(Ut.grad.numpy(), g["dU"], rtol=1e-6, atol=1e-12)
