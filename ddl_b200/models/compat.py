"""API-compatible ``Model`` for code written against the reference's ``from model import Model``.

The reference class (mnist_sync/model/model.py:5-112) exposes the dataset splits, the variable
list and three graph ops that callers run through ``self.sess``.  This class exposes the same
attribute names on top of ddl_b200: ``x_train / y_train / x_test / y_test`` (labels one-hot, as
``pd.get_dummies`` produced), ``var_bucket / var_size / var_shape``, and methods standing in for
``sess.run(self.grads)``, ``sess.run(self.accuracy)`` and ``sess.run(self.train_step)``.
"""
from __future__ import annotations

from typing import List, Optional

import torch

from ..config import Config
from ..data import load_dataset
from ..optim import FlatOptimizer
from ..parallel.arena import make_arena
from . import oracle
from .cnn import Model as _Engine
from .spec import make_spec


class Model:
    def __init__(self, cfg: Optional[Config] = None, device: Optional[str] = None):
        self.cfg = cfg or Config()
        dev = torch.device(device) if device else \
            torch.device("cuda" if (self.cfg.device != "cpu" and torch.cuda.is_available()) else "cpu")
        data = load_dataset(self.cfg.data_path, self.cfg.synthetic, self.cfg.synthetic_train,
                            self.cfg.synthetic_test, self.cfg.seed)
        self._data = data
        self.x_train, self.x_test = data.x_train, data.x_test
        self.y_train = torch.nn.functional.one_hot(data.y_train.long(), 10).float()   # model.py:13
        self.y_test = torch.nn.functional.one_hot(data.y_test.long(), 10).float()     # model.py:14
        self.spec = make_spec(self.cfg.model, self.cfg.width_mult)
        self._params = make_arena(self.spec, "none", 1, dev)
        self._grads = self._params.like()
        self._params.load_vars(oracle.init_params(self.spec, self.cfg.seed))
        self._engine = _Engine(self.spec, self._params, self._grads, self.cfg)
        self._opt = FlatOptimizer(self._params.flat, "adam", self.cfg.lr)              # model.py:93
        self.var_bucket: List[torch.Tensor] = self._params.views()                     # model.py:96
        self.var_size = len(self.var_bucket)                                           # model.py:97
        self.var_shape = [tuple(v.shape) for v in self.var_bucket]                     # model.py:98 (Q9: tuples)
        self.device = dev
        self._step = 0

    def _labels(self, y: torch.Tensor) -> torch.Tensor:
        return (y.argmax(1) if y.dim() == 2 else y).to(torch.int32)

    def compute_gradients(self, x_batch, y_batch, keep_prob: float = 0.5) -> List[torch.Tensor]:
        """``sess.run(self.grads, feed_dict={x, y_, keep_prob})`` -> 14 gradients in v0..v13 order."""
        x = torch.as_tensor(x_batch, dtype=torch.float32).to(self.device)
        y = self._labels(torch.as_tensor(y_batch)).to(self.device)
        self.last_loss = self._engine.compute_grads(x, y, self._step, keep_prob)
        self._step += 1
        return self._grads.views()

    def accuracy(self, x, y) -> float:
        """``sess.run(self.accuracy, feed_dict={..., keep_prob: 1.0})`` (model.py:104-105)."""
        x = torch.as_tensor(x, dtype=torch.float32).to(self.device)
        return self._engine.accuracy(x, self._labels(torch.as_tensor(y)).to(self.device))

    def train_step(self, x_batch, y_batch, keep_prob: float = 0.5) -> float:
        """``sess.run(self.train_step)``: gradients + local Adam (model.py:106)."""
        self.compute_gradients(x_batch, y_batch, keep_prob)
        self._opt.step(self._params.flat, self._grads.flat)
        return float(self.last_loss)

    def assign(self, values) -> None:
        """``tf.assign`` of all variables (mnist_sync/worker.py:56,69)."""
        self._params.load_vars([torch.as_tensor(v) for v in values])
