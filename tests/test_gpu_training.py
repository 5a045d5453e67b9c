"""GPU: same-seed training A/B.  The unmodified fit loops of vgan_b200.vgan are run twice on the
same device with the same seeds — once with the CUDA loss op, once with a torch-eager
restatement of the reference loss (Mmd_loss_constrained.py:24-26,42-50; test oracle) — and the
loss curves and mask statistics (per-feature selection frequency, subspace distribution) must
agree (BASELINE.json: "training must produce the same mask statistics given a fixed seed")."""
import numpy as np
import pytest
import torch
from torch import nn

pytestmark = pytest.mark.gpu


class EagerRBF(nn.Module):
    def __init__(self):
        super().__init__()
        self.bandwidth_multipliers = (2.0 ** (torch.arange(5) - 2)).cuda()
        self.bandwidth = None

    def forward(self, Z):
        L2 = torch.cdist(Z, Z, compute_mode='donot_use_mm_for_euclid_dist') ** 2
        if self.bandwidth is None:
            n = L2.shape[0]
            self.bandwidth = L2.data.sum() / (n ** 2 - n)
        return torch.exp(-L2[None, ...] / (self.bandwidth * self.bandwidth_multipliers)[:, None, None]).sum(dim=0)


class EagerLoss(nn.Module):
    def __init__(self, weight, kernel=None):
        super().__init__()
        self.kernel = EagerRBF() if kernel is None else kernel
        self.weight = weight

    def forward(self, X, Y, U):
        K = self.kernel(torch.vstack([X, Y]))
        self.bandwidth = self.kernel.bandwidth
        n = X.shape[0]
        pen = self.weight * torch.mean(1.0 - torch.max(U, 0).values)
        return K[:n, :n].mean() - 2 * K[:n, n:].mean() + K[n:, n:].mean() + pen

    def forward_masked(self, X, U):
        return self.forward(X, U * X, U)


def _data(n=1000, d=20, seed=0):
    rs = np.random.RandomState(seed)
    A = rs.randn(d, d) * 0.3 + np.eye(d)
    return (rs.randn(n, d) @ A).astype(np.float32)


def _fit(cls, loss_cls, X, monkeypatch, **kw):
    import vgan_b200.vgan as V
    from vgan_b200 import RBF
    if loss_cls is None:
        monkeypatch.setattr(V, "MMDLossConstrained", lambda weight: V.__dict__["_orig_loss"](weight, kernel=RBF()))
    else:
        monkeypatch.setattr(V, "MMDLossConstrained", lambda weight: loss_cls(weight))
    model = cls(**kw)
    model.fit(X)
    return model


@pytest.fixture(autouse=True)
def _keep_orig(monkeypatch):
    import vgan_b200.vgan as V
    from vgan_b200.loss import MMDLossConstrained
    monkeypatch.setitem(V.__dict__, "_orig_loss", MMDLossConstrained)


def test_no_kl_training_matches_eager_reference(monkeypatch, capsys):
    from vgan_b200 import VGAN_no_kl
    X = _data()
    kw = dict(batch_size=500, epochs=12, lr=0.01, seed=777)
    ours = _fit(VGAN_no_kl, None, X, monkeypatch, **kw)
    ref = _fit(VGAN_no_kl, EagerLoss, X, monkeypatch, **kw)
    a, b = np.array(ours.train_history["generator_loss"]), np.array(ref.train_history["generator_loss"])
    np.testing.assert_allclose(a, b, rtol=2e-3, atol=2e-4)
    fa = ours.generate_subspaces(500).float().mean(0).cpu().numpy()
    fb = ref.generate_subspaces(500).float().mean(0).cpu().numpy()
    assert np.abs(fa - fb).max() <= 0.02
    ours.approx_subspace_dist(500)
    ref.approx_subspace_dist(500)
    assert abs(ours.proba.max() - ref.proba.max()) <= 0.03
    assert abs(float(ours.bandwidth) - float(ref.bandwidth)) <= 1e-4 * float(ref.bandwidth)


def test_kl_training_matches_eager_reference(monkeypatch, capsys):
    from vgan_b200 import VGAN
    X = _data(1000, 32, seed=1)
    kw = dict(batch_size=500, epochs=8)
    ours = _fit(VGAN, None, X, monkeypatch, **kw)
    ref = _fit(VGAN, EagerLoss, X, monkeypatch, **kw)
    da, db = np.array(ours.train_history["detector_loss"]), np.array(ref.train_history["detector_loss"])
    np.testing.assert_allclose(da, db, rtol=2e-3, atol=1e-4)
    # the generator is detached in this class (reference quirk): its parameters never get a gradient
    assert all(p.grad is None for p in ours.generator.parameters())
    fa = ours.generate_subspaces(200).float().mean(0).cpu().numpy()
    fb = ref.generate_subspaces(200).float().mean(0).cpu().numpy()
    assert np.abs(fa - fb).max() <= 0.02


def test_check_if_myopic_runs(monkeypatch, capsys):
    from vgan_b200 import VGAN_no_kl
    X = _data(800, 16, seed=2)
    m = _fit(VGAN_no_kl, None, X, monkeypatch, batch_size=400, epochs=3, lr=0.01)
    df = m.check_if_myopic(X, bandwidth=0.01, count=300)
    assert list(df.columns) == [0.01, "recommended bandwidth"]
    assert ((df.values >= 0) & (df.values <= 1)).all()


def test_graph_replay_matches_eager_steps(monkeypatch, capsys):
    """The CUDA-graph replayed step (after 3 eager steps) and the all-eager loop train identically: same seeds, same
    RNG consumption, same kernels — only fp32 atomics order may differ."""
    from vgan_b200 import VGAN_no_kl
    X = _data()
    kw = dict(batch_size=250, epochs=6, lr=0.01, seed=777)
    graphed = _fit(VGAN_no_kl, None, X, monkeypatch, **kw)
    assert graphed.used_cuda_graph
    monkeypatch.setattr(VGAN_no_kl, "cuda_graph", False)
    eager = _fit(VGAN_no_kl, None, X, monkeypatch, **kw)
    assert not eager.used_cuda_graph
    a, b = np.array(graphed.train_history["generator_loss"]), np.array(eager.train_history["generator_loss"])
    np.testing.assert_allclose(a, b, rtol=1e-4, atol=1e-6)
    fa = graphed.generate_subspaces(500).float().mean(0).cpu().numpy()
    fb = eager.generate_subspaces(500).float().mean(0).cpu().numpy()
    assert np.abs(fa - fb).max() <= 0.01


def test_kl_graph_replay_matches_eager_steps(monkeypatch, capsys):
    from vgan_b200 import VGAN
    X = _data(1000, 32, seed=1)
    kw = dict(batch_size=250, epochs=8)
    graphed = _fit(VGAN, None, X, monkeypatch, **kw)
    assert graphed.used_cuda_graph
    monkeypatch.setattr(VGAN, "cuda_graph", False)
    eager = _fit(VGAN, None, X, monkeypatch, **kw)
    assert not eager.used_cuda_graph
    for key in ("detector_loss", "generator_loss"):
        a, b = np.array(graphed.train_history[key]), np.array(eager.train_history[key])
        np.testing.assert_allclose(a, b, rtol=2e-4, atol=1e-6, equal_nan=True)


def test_graph_replay_wide_data_uses_cluster_kernels(monkeypatch, capsys):
    """D > 256: the captured step contains the pair kernel and the S_xx kernel on their internal streams (fork/join
    inside the capture); graph and eager training agree."""
    from vgan_b200 import VGAN_no_kl
    X = _data(768, 320, seed=4)
    kw = dict(batch_size=256, epochs=4, lr=0.01, seed=777)
    graphed = _fit(VGAN_no_kl, None, X, monkeypatch, **kw)
    assert graphed.used_cuda_graph
    monkeypatch.setattr(VGAN_no_kl, "cuda_graph", False)
    eager = _fit(VGAN_no_kl, None, X, monkeypatch, **kw)
    a, b = np.array(graphed.train_history["generator_loss"]), np.array(eager.train_history["generator_loss"])
    np.testing.assert_allclose(a, b, rtol=1e-4, atol=1e-6)
