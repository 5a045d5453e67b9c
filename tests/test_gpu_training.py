"""GPU: same-seed training A/B.  The unmodified fit loops of vgan_b200.vgan are run twice on the
same device with the same seeds — once with the CUDA loss op, once with the pinned torch-eager
restatement of the reference loss (oracle/torch_eager.py; Mmd_loss_constrained.py:24-26,42-50) — and the
loss curves and mask statistics (per-feature selection frequency, subspace distribution) must
agree (BASELINE.json: "training must produce the same mask statistics given a fixed seed")."""
import numpy as np
import pytest
import torch

from oracle import torch_eager

pytestmark = pytest.mark.gpu


class EagerLoss(torch_eager.MMDLossConstrained):
    """The pinned torch restatement of the reference module (oracle/torch_eager.py, checked against the golden
    vectors in tests/test_oracle.py) with the one extra call the re-plumbed no-KL step makes."""

    def __init__(self, weight):
        super().__init__(weight, kernel=torch_eager.RBF(device="cuda"), device="cuda")

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


@pytest.mark.parametrize("d,bs", [(96, 512), (320, 256)])
def test_no_kl_training_default_bf16_path_matches_eager_reference(monkeypatch, capsys, d, bs):
    """The DEFAULT precision at realistic widths (D >= 32: bf16 operands on the tensor cores; D = 320: the multi-slice
    CTA-pair kernel) inside the training loop, against the eager reference loss: loss curves and mask statistics."""
    from vgan_b200 import VGAN_no_kl
    X = _data(1024, d, seed=3)
    kw = dict(batch_size=bs, epochs=8, lr=0.01, seed=777)
    ours = _fit(VGAN_no_kl, None, X, monkeypatch, **kw)
    ref = _fit(VGAN_no_kl, EagerLoss, X, monkeypatch, **kw)
    a, b = np.array(ours.train_history["generator_loss"]), np.array(ref.train_history["generator_loss"])
    np.testing.assert_allclose(a, b, rtol=2e-3, atol=2e-4)
    fa = ours.generate_subspaces(500).float().mean(0).cpu().numpy()
    fb = ref.generate_subspaces(500).float().mean(0).cpu().numpy()
    assert np.abs(fa - fb).max() <= 0.03


def test_kl_detector_encoder_stays_frozen_after_first_generator_phase(monkeypatch, capsys):
    """Reference quirk (src/vgan.py:266-267, 319-320): the generator phase freezes every detector parameter and the
    detector phase re-enables the decoder only, so the encoder is trained in the first detector phase and never again."""
    from vgan_b200 import VGAN
    X = _data(1000, 32, seed=1)
    m = _fit(VGAN, None, X, monkeypatch, batch_size=500, epochs=8)      # D, G x5, D, G
    assert all(not p.requires_grad for p in m.detector.encoder.parameters())
    m1 = _fit(VGAN, None, X, monkeypatch, batch_size=500, epochs=1)     # after the first detector phase
    m7 = _fit(VGAN, None, X, monkeypatch, batch_size=500, epochs=7)     # after the second detector phase
    for a, b in zip(m1.detector.encoder.parameters(), m7.detector.encoder.parameters()):
        assert float((a - b).abs().max()) <= 2e-6      # equal up to the atomics' summation order in epoch 0
    # the decoder did train in the second detector phase (two Adadelta steps at lr 0.007: a few 1e-5 per weight)
    assert any(float((a.detach() - b.detach()).abs().max()) > 1e-5
               for a, b in zip(m1.detector.decoder.parameters(), m7.detector.decoder.parameters()))


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


def test_no_kl_fit_within_reference_band(monkeypatch, capsys):
    """vgan_b200.VGAN_no_kl.fit against the band spanned by the REFERENCE's own loop (tests/golden/fit_no_kl.npz, made by
    oracle/make_fit_fixtures.py from /root/reference/src/vgan.py:546-637 on CPU for five seeds).  Bit equality is
    impossible (the reference draws batches from its DataLoader's RNG and noise from the CPU generator, this build from
    device generators), and the reference itself is strongly seed dependent — the epoch loss is dominated by the
    penalty, 10 x the fraction of features no sampled mask selects, in steps of 0.24 — so what is held is: every epoch
    loss inside the band the reference runs span (+- two penalty steps), the seed-averaged loss level and the seed-averaged number of
    selected features close to the reference's, and the same frozen bandwidth per seed (it depends on the first batch
    only through which rows it holds: a few per cent)."""
    import os
    from vgan_b200 import VGAN_no_kl
    from oracle.make_fit_fixtures import BATCH, EPOCHS, LR, SEEDS, fit_data
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "fit_no_kl.npz"))
    X = fit_data()
    losses, nsel, bws = [], [], []
    for seed in SEEDS:
        m = _fit(VGAN_no_kl, None, X, monkeypatch, batch_size=BATCH, epochs=EPOCHS, lr=LR, seed=int(seed))
        losses.append(np.array(m.train_history["generator_loss"]))
        nsel.append(float(m.generate_subspaces(500).float().sum(1).mean()))
        bws.append(float(m.bandwidth))
    L = np.stack(losses)
    # the band over all seeds and epochs (five reference runs do not pin a band PER epoch: their levels jump by whole
    # penalty steps between epochs), widened by two penalty steps of 0.24
    lo, hi = float(g["loss"].min()), float(g["loss"].max())
    assert L.min() >= lo - 0.5 and L.max() <= hi + 0.5, (L.min(), L.max(), lo, hi)
    assert abs(L.mean() - g["loss"].mean()) <= 0.8, (L.mean(), g["loss"].mean())
    ref_nsel = float(g["freq"].sum(1).mean())
    assert abs(np.mean(nsel) - ref_nsel) <= 3.0, (nsel, ref_nsel)
    assert np.abs(np.array(bws) / g["bandwidth"].mean() - 1.0).max() <= 0.1, (bws, g["bandwidth"])
