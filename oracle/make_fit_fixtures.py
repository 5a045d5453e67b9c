"""Generate tests/golden/fit_no_kl.npz by running the REFERENCE's own training loop (src/vgan.py, VGAN_no_kl.fit) on CPU.

TEST INFRASTRUCTURE ONLY.  Run in the authoring container (needs /root/reference):

    python oracle/make_fit_fixtures.py

`import src.vgan` needs torch_two_sample and matplotlib, neither installed nor used by `fit`: two
empty stub modules are registered first.  For a few seeds the script fits VGAN_no_kl(epochs=20, batch_size=500,
lr=0.01) to seeded N=1000, D=20 float32 data (a correlated Gaussian: some features redundant) and stores per epoch
the mean batch loss and, after training, the per-feature selection frequency of 500 sampled subspaces and the
probability of the most frequent subspace.  The batch order comes from the reference's DataLoader RNG, ours from a
device permutation, so a run of vgan_b200.VGAN_no_kl can only be held to the BAND these runs span (plus a margin), not
to their digits: tests/test_gpu_training.py::test_no_kl_fit_within_reference_band.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(os.path.dirname(HERE), "tests", "golden")
SEEDS = (0, 1, 2, 3, 4)
EPOCHS, BATCH, LR = 20, 500, 0.01


def fit_data(n=1000, d=20, seed=0) -> np.ndarray:
    """Correlated Gaussian: a random mixing matrix close to the identity (same recipe as tests/test_gpu_training.py)."""
    rs = np.random.RandomState(seed)
    A = rs.randn(d, d) * 0.3 + np.eye(d)
    return (rs.randn(n, d) @ A).astype(np.float32)


def load_reference_vgan():
    for name in ("torch_two_sample", "matplotlib", "matplotlib.pyplot"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.path.insert(0, "/root/reference")
    import src.vgan as ref_vgan  # type: ignore
    return ref_vgan


def run_reference(seed: int):
    ref = load_reference_vgan()
    # the reference's default-argument RBF() is shared process-wide and freezes its bandwidth on the first call
    # (Mmd_loss_constrained.py:16-22,35): give every run a fresh one, as a fresh process would
    from src.models.Mmd_loss_constrained import RBF  # type: ignore
    ref.MMDLossConstrained.__init__.__defaults__ = (RBF(),)
    X = fit_data()
    model = ref.VGAN_no_kl(batch_size=BATCH, epochs=EPOCHS, lr=LR, seed=seed)
    model.fit(X)
    u = model.generate_subspaces(500)
    freq = u.float().mean(0).cpu().numpy()
    model.approx_subspace_dist(500)
    return dict(loss=np.array(model.train_history["generator_loss"], dtype=np.float64), freq=freq,
                top_proba=float(np.max(model.proba)), bandwidth=float(model.bandwidth))


def main():
    runs = [run_reference(s) for s in SEEDS]
    os.makedirs(GOLDEN, exist_ok=True)
    np.savez_compressed(os.path.join(GOLDEN, "fit_no_kl.npz"), seeds=np.array(SEEDS),
                        loss=np.stack([r["loss"] for r in runs]), freq=np.stack([r["freq"] for r in runs]),
                        top_proba=np.array([r["top_proba"] for r in runs]),
                        bandwidth=np.array([r["bandwidth"] for r in runs]),
                        epochs=EPOCHS, batch=BATCH, lr=LR)
    L = np.stack([r["loss"] for r in runs])
    print("per-epoch loss min/max over seeds:")
    for e in range(EPOCHS):
        print(f"  {e:2d}  {L[:, e].min():.5f} .. {L[:, e].max():.5f}")
    F = np.stack([r["freq"] for r in runs])
    print("feature frequency mean over seeds:", np.round(F.mean(0), 2))
    print("feature frequency spread (max-min):", np.round(F.max(0) - F.min(0), 2))
    print("top subspace proba:", [round(r["top_proba"], 3) for r in runs], "bandwidth", [round(r["bandwidth"], 4) for r in runs])


if __name__ == "__main__":
    main()
