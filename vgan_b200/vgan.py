"""Host mirror of /root/reference/src/vgan.py: ``VGAN`` (kernel learning, :20-431) and
``VGAN_no_kl`` (:434-708) with the same constructor and method signatures, built around the
B200 loss op.  What is re-plumbed and what is kept on purpose:

* the loss module is ``vgan_b200.MMDLossConstrained`` (one fused kernel pass per step); the
  no-KL step calls ``forward_masked`` so ``U*X`` and ``dL/dU`` never exist as torch ops;
* the data set is moved to the device once and batches are drawn with a device permutation
  (same ``drop_last`` / shuffle semantics as the per-epoch ``DataLoader``, vgan.py:221-227,578-584),
  and the epoch loss is accumulated on the device and read once per epoch instead of a
  ``.to('cpu')`` sync per step (:288,:328,:620);
* kept: seeds (``VGAN`` forces 777, :48), Adadelta + weight decay, the 1-detector/5-generator
  epoch schedule (:251-332), ``weight=temperature`` resp. ``weight=10`` (:214,:571), the
  ``Variable(generator(noise))`` detach in the generator phase of ``VGAN`` which leaves the
  generator without gradients (:307-310), the encoder that stays frozen after the first
  generator phase (:266-267 re-enable the decoder only, :319-320 freeze the whole detector), re-seeding in
  ``generate_subspaces`` (:364-366), the unique-rows subspace distribution (:372-382) and the
  ``check_if_myopic`` protocol (:384-431) with ``alphas=[bandwidth]`` passed as-is.
The generator/detector MLPs stay in torch (models.py)."""
from __future__ import annotations

import os
from collections import defaultdict
from pathlib import Path
from typing import Union

import numpy as np
import pandas as pd
import torch

from .loss import MMDLossConstrained
from .models import Decoder, Detector, Encoder, Generator_big
from .statistic import MMDStatistic
from .subspace import unique_subspaces as unique_subspaces_device


def _device():
    return torch.device('cuda:0' if torch.cuda.is_available() else 'cpu')


class _GraphedStep:
    """Runs ``fn(batch)`` eagerly for ``warmup`` calls, then captures it once into a CUDA graph on a static batch
    buffer (nothing executes during capture) and replays it for every later batch.  The steps are launch-bound at
    the reference's batch sizes (N = 500); a replay costs one launch."""

    def __init__(self, fn, enabled: bool, warmup: int = 3):
        self.fn, self.enabled, self.warmup = fn, enabled, warmup
        self.graph, self.static, self.calls, self.key = None, None, 0, None

    def __call__(self, batch, key=None):
        """``key``: anything the captured work depends on besides the batch (e.g. which parameters are trainable);
        a change drops the graph and starts over with eager warm-up steps."""
        if key != self.key:
            self.graph, self.static, self.calls, self.key = None, None, 0, key
        if self.graph is not None:
            self.static.copy_(batch)
            self.graph.replay()
            return
        self.fn(batch)
        self.calls += 1
        if self.enabled and self.calls == self.warmup:
            self.static = torch.empty_like(batch)
            graph = torch.cuda.CUDAGraph()
            torch.cuda.synchronize()
            with torch.cuda.graph(graph):
                self.fn(self.static)
            self.graph = graph


class _SubspaceGAN:
    """Everything the two reference classes share."""

    def _seed_all(self):
        torch.manual_seed(self.seed)
        if torch.cuda.is_available():
            torch.cuda.manual_seed(self.seed)

    def _to_device(self, X):
        # the reference casts each batch to float32 on cuda (vgan.py:256-263, 600-607)
        return torch.as_tensor(np.asarray(X)).to(torch.float32).to(self.device).reshape(len(X), -1)

    def _batches(self, Xd):
        """drop_last, shuffled batches of the resident data set (one device permutation per epoch)."""
        n = Xd.shape[0]
        perm = torch.randperm(n, device=Xd.device)
        for b in range(n // self.batch_size):
            yield Xd[perm[b * self.batch_size:(b + 1) * self.batch_size]]

    cuda_graph = True            # replay the training step as a CUDA graph (VGAN_B200_NO_GRAPH=1 or False: eager)
    _graph_warmup_steps = 3

    def _use_graph(self) -> bool:
        return bool(self.cuda_graph and self.device.type == 'cuda' and not os.environ.get('VGAN_B200_NO_GRAPH'))

    def _save_run(self, generator, also_detector_slot: bool):
        if self.path_to_directory is None:
            return
        root = Path(self.path_to_directory)
        (root / 'models').mkdir(parents=True, exist_ok=True)
        n_files = len(os.listdir(root / 'models'))
        run_number = n_files // 2 if also_detector_slot else n_files
        torch.save(generator.state_dict(), root / 'models' / f'generator_{run_number}.pt')
        if also_detector_slot:   # vgan.py:348-349 stores the generator under the detector name too
            torch.save(generator.state_dict(), root / 'models' / f'detector_{run_number}.pt')
        self.model_snapshot(root, run_number, show=True)

    def model_snapshot(self, path_to_directory=None, run_number=0, show=False):
        root = Path(self.path_to_directory if path_to_directory is None else path_to_directory)
        (root / 'train_history').mkdir(parents=True, exist_ok=True)
        for key in ('generator_loss', 'detector_loss'):
            if key in self.train_history:
                pd.DataFrame(self.train_history[key]).to_csv(
                    root / 'train_history' / f'{key}_{run_number}.csv', header=False, index=False)
        new = pd.DataFrame(self.get_params(), [run_number])
        if (root / 'params.csv').is_file():
            old = pd.read_csv(root / 'params.csv', index_col=0)
            old = old.reindex(old.index.union(new.index))
            old.update(new)
            new = old
        new.to_csv(root / 'params.csv')
        try:                               # the PDF loss plot (vgan.py:80-99) needs matplotlib
            import matplotlib
            matplotlib.use('Agg')
            import matplotlib.pyplot as plt
            fig, ax = plt.subplots()
            for key in ('generator_loss', 'detector_loss'):
                if key in self.train_history:
                    ax.plot(self.train_history[key], label=key)
            ax.legend()
            fig.savefig(root / 'train_history' / f'losses_{run_number}.pdf')
            plt.close(fig)
        except ImportError:
            pass

    def load_models(self, path_to_generator, ndims, device: str = None):
        device = self.device if device is None else device
        self._latent_size = max(int(ndims / 16), 1)
        self.generator = Generator_big(img_size=ndims, latent_size=self._latent_size).to(device)
        self.generator.load_state_dict(torch.load(path_to_generator, map_location=device))
        self.generator.eval()
        self.generator_optimizer = f'Loaded Model from {path_to_generator} with {ndims} dimensions in the latent space'

    def generate_subspaces(self, nsubs) -> torch.Tensor:
        noise = torch.Tensor(nsubs, self._latent_size).to('cpu')
        if self.seed is not None:
            torch.manual_seed(self.seed)      # every call returns the same masks (vgan.py:364-366)
        noise.normal_()
        u = self.generator(noise.to(self.device))
        return torch.greater_equal(u, 1 / u.shape[1])

    def approx_subspace_dist(self, subspace_count=500, add_leftover_features=False):
        u = self.generate_subspaces(subspace_count)
        # unique rows + multiplicities on the device, in np.unique(axis=0) order (subspace.py)
        uniq_d, counts_d = unique_subspaces_device(u)
        unique_subspaces, proba = uniq_d.cpu().numpy(), counts_d.cpu().numpy()
        unused = unique_subspaces.sum(axis=0) < 1
        if unused.sum() != 0 and add_leftover_features:
            unique_subspaces = np.append(unique_subspaces, [unused], axis=0)
            proba = np.append(proba / proba.sum(), 1)
        self.subspaces = unique_subspaces
        self.proba = proba / proba.sum()

    def check_if_myopic(self, x_data: np.array, bandwidth: Union[float, np.array] = 0.01, count=500) -> pd.DataFrame:
        assert count <= x_data.shape[0], "Selected 'count' is greater than the number of samples in the dataset"
        x_data = np.asarray(x_data, dtype=np.float64)
        norms = np.sqrt((x_data ** 2).sum(axis=0, keepdims=True))
        norms[norms == 0] = 1.0
        x_data = x_data / norms                                  # sklearn normalize(axis=0)
        x_sample = torch.Tensor(pd.DataFrame(x_data).sample(count).to_numpy()).to(self.device)
        u = self.generate_subspaces(count)
        ux_sample = u * x_sample + torch.mean(x_sample, dim=0) * (~u)
        bandwidth = [bandwidth] if isinstance(bandwidth, float) else list(bandwidth)
        if not hasattr(self, 'bandwidth'):
            probe = MMDLossConstrained(0)
            with torch.no_grad():
                probe.forward(x_sample, ux_sample, u * 1)
            self.bandwidth = probe.bandwidth
        bandwidth.sort()
        results = []
        for bw in bandwidth + [self.bandwidth.item()]:
            mmd = MMDStatistic(count, count)
            _, distances = mmd(x_sample, ux_sample, alphas=[bw], ret_matrix=True)
            results.append(mmd.pval(distances))
        return pd.DataFrame([results], columns=bandwidth + ["recommended bandwidth"], index=["p-val"])


class VGAN(_SubspaceGAN):
    '''
    VGAN with kernel learning (reference class `VGAN`, src/vgan.py:20-431): the MMD is taken
    between detector encodings of the batch and of the masked batch.
    '''

    def __init__(self, batch_size=500, temperature=0, epochs=2000, lr_G=0.007, lr_D=0.007, iternum_d=1,
                 iternum_g=5, momentum=0.99, seed=777, weight_decay=0.04, path_to_directory=None):
        self.storage = locals()
        self.train_history = defaultdict(list)
        self.batch_size = batch_size
        self.temperature = temperature
        self.epochs = epochs
        self.lr_G = lr_G
        self.lr_D = lr_D
        self.iternum_d = iternum_d
        self.iternum_g = iternum_g
        self.momentum = momentum
        self.weight_decay = weight_decay
        self.path_to_directory = path_to_directory
        self.generator_optimizer = None
        self.device = _device()
        self.seed = 777                     # the reference overrides the argument (vgan.py:48)

    def get_params(self) -> dict:
        return {'batch size': self.batch_size, 'epochs': self.epochs, 'lr_g': self.lr_G,
                'momentum': self.momentum, 'weight decay': self.weight_decay,
                'batch_size': self.batch_size, 'seed': self.seed,
                'generator optimizer': self.generator_optimizer}

    def get_the_networks(self, ndims: int, latent_size: int, device: str = None) -> tuple:
        device = self.device if device is None else device
        generator = Generator_big(img_size=ndims, latent_size=latent_size).to(device)
        detector = Detector(latent_size, ndims, Encoder, Decoder).to(device)
        return generator, detector

    @staticmethod
    def _weights_init(m):
        if m.__class__.__name__.find('Linear') != -1:     # vgan.py:69-78
            m.weight.data.normal_(0.0, 0.1)
            m.bias.data.fill_(0)

    def fit(self, X):
        self._seed_all()
        self._latent_size = latent_size = max(int(X.shape[1] / 16), 1)
        ndims = X.shape[1]
        self.batch_size = min(self.batch_size, X.shape[0])
        generator, detector = self.get_the_networks(ndims, latent_size, device=self.device)
        generator.apply(self._weights_init)
        detector.apply(self._weights_init)
        use_graph = self._use_graph()
        gen_optimizer = torch.optim.Adadelta(generator.parameters(), lr=self.lr_G, weight_decay=self.weight_decay,
                                             capturable=use_graph)
        det_optimizer = torch.optim.Adadelta(detector.parameters(), lr=self.lr_D, weight_decay=self.weight_decay,
                                             capturable=use_graph)
        self.generator_optimizer = gen_optimizer.__class__.__name__
        self.detector_optimizer = det_optimizer.__class__.__name__
        loss_function = MMDLossConstrained(weight=self.temperature)
        Xd = self._to_device(X)
        batch_number = Xd.shape[0] // self.batch_size
        noise = torch.empty(self.batch_size, latent_size, device=self.device)
        l2 = lambda a, b: (a - b).pow(2).mean()     # noqa: E731
        acc = torch.zeros((), device=self.device)

        def detector_step(batch):
            # maximise MMD(enc, penc) - 0.1 (reconstruction terms)  (:251-291).  Only the DECODER is re-enabled here
            # (:266-267): the generator phase freezes every detector parameter (:319-320), so from the second detector
            # phase on the encoder stays frozen and only the reconstruction terms train the decoder.  Kept.
            for p in detector.decoder.parameters():
                p.requires_grad = True
            enc, dec = detector(batch)
            with torch.no_grad():
                U = generator(noise.normal_()).clone().detach()
            masked = U * batch
            penc, pdec = detector(masked)
            det_optimizer.zero_grad()
            loss_D = -(loss_function(enc, penc, U) - .1 * l2(batch, dec) - .1 * l2(masked, pdec))
            loss_D.backward()
            det_optimizer.step()
            acc.add_(loss_D.detach() / batch_number)

        def generator_step(batch):
            # (:293-332)  `Variable(generator(noise))` in the reference is a detached leaf: the MMD gradient stops
            # at U and the generator gets none.  Kept.
            # The detector's (unused) gradients of this phase are dropped first so that a captured step allocates
            # and frees them inside its own graph instead of accumulating into another graph's buffers.
            det_optimizer.zero_grad()
            enc, _ = detector(batch)
            U = generator(noise.normal_()).detach().requires_grad_(True)
            penc, _ = detector(U * batch)
            for p in detector.parameters():
                p.requires_grad = False
            gen_optimizer.zero_grad()
            loss_G = loss_function(enc, penc, U)
            loss_G.backward()
            gen_optimizer.step()
            acc.add_(loss_G.detach() / batch_number)

        run_detector = _GraphedStep(detector_step, use_graph, self._graph_warmup_steps)
        run_generator = _GraphedStep(generator_step, use_graph, self._graph_warmup_steps)
        iternum_d = iternum_g = 1
        detector_loss = generator_loss = np.nan
        for epoch in range(self.epochs):
            print(f'\rEpoch {epoch} of {self.epochs}')
            acc.zero_()
            if iternum_d <= self.iternum_d:
                for batch in self._batches(Xd):
                    # the captured step depends on which parameters are trainable (all of them in the first detector
                    # phase, the decoder only afterwards): re-captured when that changes
                    run_detector(batch, key=tuple(p.requires_grad for p in detector.parameters()))
                self.bandwidth = loss_function.bandwidth
                detector_loss = float(acc)
                iternum_d += 1
                iternum_g = 1
            elif iternum_g <= self.iternum_g:
                for batch in self._batches(Xd):
                    run_generator(batch)
                self.bandwidth = loss_function.bandwidth
                generator_loss = float(acc)
                iternum_g += 1
                if iternum_g > self.iternum_g:
                    iternum_d = 1
            print(f"Average loss in the epoch Generator: {generator_loss}")
            print(f"Average loss in the epoch Detector: {detector_loss}")
            self.train_history["generator_loss"].append(generator_loss)
            self.train_history["detector_loss"].append(detector_loss)
        self.used_cuda_graph = run_detector.graph is not None or run_generator.graph is not None
        self._save_run(generator, also_detector_slot=True)
        self.generator = generator
        self.detector = detector


class VGAN_no_kl(_SubspaceGAN):
    '''
    VGAN without kernel learning (reference class `VGAN_no_kl`, src/vgan.py:434-708): the MMD is
    taken between the raw batch and the masked batch u*x — the path BASELINE.json's north star names.
    '''

    def __init__(self, batch_size=500, epochs=2000, lr=0.007, momentum=0.99, seed=777, weight_decay=0.04,
                 path_to_directory=None):
        self.storage = locals()
        self.train_history = defaultdict(list)
        self.batch_size = batch_size
        self.epochs = epochs
        self.lr = lr
        self.momentum = momentum
        self.seed = seed
        self.weight_decay = weight_decay
        self.path_to_directory = path_to_directory
        self.generator_optimizer = None
        self.device = _device()

    def get_params(self) -> dict:
        return {'batch size': self.batch_size, 'epochs': self.epochs, 'lr_g': self.lr,
                'momentum': self.momentum, 'weight decay': self.weight_decay,
                'batch_size': self.batch_size, 'seed': self.seed,
                'generator optimizer': self.generator_optimizer}

    def get_the_networks(self, ndims: int, latent_size: int, device: str = None) -> Generator_big:
        device = self.device if device is None else device
        return Generator_big(img_size=ndims, latent_size=latent_size).to(device)

    def fit(self, X):
        self._seed_all()
        self._latent_size = latent_size = max(int(X.shape[1] / 16), 1)
        ndims = X.shape[1]
        self.batch_size = min(self.batch_size, X.shape[0])
        generator = self.get_the_networks(ndims, latent_size, device=self.device)
        use_graph = self._use_graph()
        optimizer = torch.optim.Adadelta(generator.parameters(), lr=self.lr, weight_decay=self.weight_decay,
                                         capturable=use_graph)
        self.generator_optimizer = optimizer.__class__.__name__
        loss_function = MMDLossConstrained(weight=10)
        Xd = self._to_device(X)
        batch_number = Xd.shape[0] // self.batch_size
        noise = torch.empty(self.batch_size, latent_size, device=self.device)
        acc = torch.zeros((), device=self.device)

        def step(batch):
            noise.normal_()
            optimizer.zero_grad(set_to_none=True)
            fake_subspaces = generator(noise)
            batch_loss = loss_function.forward_masked(batch, fake_subspaces)   # loss(x, u*x, u)
            batch_loss.backward()
            optimizer.step()
            acc.add_(batch_loss.detach() / batch_number)

        # noise -> generator -> prologue -> fused MMD kernel -> backward -> Adadelta: captured into a CUDA graph after
        # a few eager steps (bandwidth frozen, optimizer state and workspaces allocated), then replayed per batch
        run_step = _GraphedStep(step, use_graph, self._graph_warmup_steps)
        for epoch in range(self.epochs):
            print(f'\rEpoch {epoch} of {self.epochs}')
            acc.zero_()
            for batch in self._batches(Xd):
                run_step(batch)
            self.bandwidth = loss_function.bandwidth
            generator_loss = float(acc)                 # one host sync per epoch
            print(f"Average loss in the epoch: {generator_loss}")
            self.train_history["generator_loss"].append(generator_loss)
        self.used_cuda_graph = run_step.graph is not None
        self._save_run(generator, also_detector_slot=False)
        self.generator = generator
