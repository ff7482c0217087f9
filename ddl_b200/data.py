"""Dataset layer: the reference's ``data/mnist.pkl`` loader and a synthetic stand-in.

mnist_sync/model/model.py:8-14 loads a latin1 pickle holding ``(train, valid, test)`` where
each split is ``(x[N,784] float32, y[N] int)`` and one-hots the labels with
``pd.get_dummies``.  We keep integer labels (the fused softmax-xent kernel takes class ids;
a one-hot view is available for parity) and add a deterministic synthetic generator of the
same shape because no dataset can be downloaded here.

Hot-path note: the whole training split lives in pinned host memory; each step's slice is
copied with one async H2D per tensor (never the batched-memcpy driver entry points).
"""
from __future__ import annotations

import os
import pickle
from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np
import torch


@dataclass
class Dataset:
    x_train: torch.Tensor      # [N,784] float32 (CPU, pinned when CUDA is present)
    y_train: torch.Tensor      # [N] int32 class ids
    x_test: torch.Tensor
    y_test: torch.Tensor
    synthetic: bool

    @property
    def num_train(self) -> int:
        return int(self.x_train.shape[0])

    def one_hot_train(self) -> torch.Tensor:
        """Parity with ``pd.get_dummies(y_train)`` (model.py:13)."""
        return torch.nn.functional.one_hot(self.y_train.long(), 10).float()

    def batch_slice(self, cnt: int, batch_size: int, worker: int = 0, num_workers: int = 1,
                    replicate: bool = False) -> Tuple[int, int]:
        """[begin, end) rows of batch ``cnt`` for ``worker``.

        The reference gives every worker rows ``[bs*cnt, bs*(cnt+1))`` with no rank offset
        (mnist_sync/worker.py:27-28, quirk Q6) - that is ``replicate=True``.  The default
        shards the epoch by rank: worker w takes the w-th batch of every group of W batches.
        """
        if replicate or num_workers == 1:
            b = batch_size * cnt
        else:
            b = batch_size * (cnt * num_workers + worker)
        n = self.num_train
        b = b % max(1, n - batch_size + 1)
        return b, b + batch_size

    def batches_per_epoch(self, batch_size: int, num_workers: int = 1, replicate: bool = False) -> int:
        nb = self.num_train // batch_size          # int(N / batch_size), worker.py:59
        return nb if (replicate or num_workers == 1) else max(1, nb // num_workers)


def _pin(t: torch.Tensor) -> torch.Tensor:
    if torch.cuda.is_available():
        try:
            return t.pin_memory()
        except RuntimeError:
            return t
    return t


def synthetic_mnist(n_train: int = 6000, n_test: int = 1000, seed: int = 1234) -> Dataset:
    """Class-conditional blobs in 784-d: learnable (accuracy rises within tens of steps) yet
    cheap.  Each class has a fixed random template; samples are template + noise, clipped to
    [0,1] like MNIST pixels."""
    g = torch.Generator().manual_seed(seed)
    templates = (torch.rand(10, 784, generator=g) < 0.15).float()

    def make(n, gen):
        y = torch.randint(0, 10, (n,), generator=gen, dtype=torch.int64)
        x = templates[y] * (0.6 + 0.4 * torch.rand(n, 784, generator=gen)) \
            + 0.25 * torch.rand(n, 784, generator=gen) * (torch.rand(n, 784, generator=gen) < 0.1)
        return x.clamp_(0, 1).float().contiguous(), y.to(torch.int32)

    xt, yt = make(n_train, g)
    xe, ye = make(n_test, g)
    return Dataset(_pin(xt), _pin(yt), _pin(xe), _pin(ye), True)


def load_mnist_pkl(path: str) -> Dataset:
    """Reads the reference's pickle: 3-tuple, latin1 (model.py:8-11)."""
    with open(path, "rb") as f:
        train_set, _, test_set = pickle.load(f, encoding="latin1")
    x_train, y_train = train_set
    x_test, y_test = test_set
    conv = lambda x: torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32))
    lab = lambda y: torch.from_numpy(np.ascontiguousarray(y, dtype=np.int32))
    return Dataset(_pin(conv(x_train)), _pin(lab(y_train)), _pin(conv(x_test)), _pin(lab(y_test)), False)


def load_dataset(path: str = "data/mnist.pkl", synthetic: bool = False, n_train: int = 6000,
                 n_test: int = 1000, seed: Optional[int] = None) -> Dataset:
    if not synthetic and os.path.exists(path):
        return load_mnist_pkl(path)
    if not synthetic:
        print("[ddl_b200] %s not found - using synthetic MNIST-shaped data" % path, flush=True)
    return synthetic_mnist(n_train, n_test, 1234 if seed is None else seed)
