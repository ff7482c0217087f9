"""Variable inventory of the reference model (the numerical contract).

Restates the 14 ``get_variable`` calls of mnist_sync/model/model.py:24-86: names
``v0..v13``, HWIO conv kernels, ``[in, out]`` FC matrices, (h, w, c) flatten order.
``width_mult`` widens every channel / hidden size by an integer factor (BASELINE.json
config 5, the ~100M-parameter communication-bound point); ``width_mult=1`` is the
reference exactly (2,656,010 parameters).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Tuple

IMAGE_HW = 28
NUM_CLASSES = 10
KSIZE = 5


@dataclass(frozen=True)
class VarSpec:
    index: int
    name: str                 # "v0" .. "v13"  (scope "mnist/", model.py:17)
    shape: Tuple[int, ...]
    kind: str                 # conv_w | fc_w | bias
    layer: str                # conv1..conv4, fc1..fc3

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n

    @property
    def nbytes(self) -> int:
        return 4 * self.numel

    def fans(self) -> Tuple[int, int]:
        """TF1 glorot fan computation (variables have no explicit initializer, model.py:24)."""
        if len(self.shape) == 1:
            return self.shape[0], self.shape[0]
        if len(self.shape) == 2:
            return self.shape[0], self.shape[1]
        rf = 1
        for s in self.shape[:-2]:
            rf *= s
        return rf * self.shape[-2], rf * self.shape[-1]


@dataclass(frozen=True)
class ModelSpec:
    kind: str                         # cnn | mlp
    width_mult: int
    conv_channels: Tuple[int, ...]    # (1, 32, 64, 128, 256) * wm
    fc_sizes: Tuple[int, ...]         # (flat, 1024, 512, 10)
    vars: Tuple[VarSpec, ...]

    @property
    def num_vars(self) -> int:
        return len(self.vars)

    @property
    def num_params(self) -> int:
        return sum(v.numel for v in self.vars)

    def spatial(self) -> List[int]:
        """Spatial size at the input of each conv layer: 28, 14, 7, 4 and the final 2."""
        out, s = [], IMAGE_HW
        for _ in range(len(self.conv_channels) - 1):
            out.append(s)
            s = (s + 1) // 2          # 2x2 stride-2 SAME pool: ceil(s/2); 7 -> 4 (model.py:53)
        out.append(s)
        return out

    def fwd_flops(self, batch: int) -> float:
        f, sp = 0.0, self.spatial()
        for i in range(len(self.conv_channels) - 1):
            f += 2.0 * batch * sp[i] * sp[i] * KSIZE * KSIZE * self.conv_channels[i] * self.conv_channels[i + 1]
        for i in range(len(self.fc_sizes) - 1):
            f += 2.0 * batch * self.fc_sizes[i] * self.fc_sizes[i + 1]
        return f

    def step_flops(self, batch: int) -> float:
        """fwd + dgrad + wgrad; conv1 has no dgrad."""
        sp = self.spatial()
        c1 = 2.0 * batch * sp[0] * sp[0] * KSIZE * KSIZE * self.conv_channels[0] * self.conv_channels[1] \
            if self.kind == "cnn" else 0.0
        return 3.0 * self.fwd_flops(batch) - c1


def make_spec(kind: str = "cnn", width_mult: int = 1) -> ModelSpec:
    wm = int(width_mult)
    if wm < 1:
        raise ValueError("width_mult must be >= 1")
    vs: List[VarSpec] = []

    def add(shape, k, layer):
        vs.append(VarSpec(len(vs), "v%d" % len(vs), tuple(shape), k, layer))

    if kind == "cnn":
        ch = (1, 32 * wm, 64 * wm, 128 * wm, 256 * wm)
        for i in range(4):                                    # model.py:24-64
            add((KSIZE, KSIZE, ch[i], ch[i + 1]), "conv_w", "conv%d" % (i + 1))
            add((ch[i + 1],), "bias", "conv%d" % (i + 1))
        flat = 2 * 2 * ch[4]                                  # model.py:67,69
        fc = (flat, 1024 * wm, 512 * wm, NUM_CLASSES)
    elif kind == "mlp":
        ch = (1,)
        fc = (IMAGE_HW * IMAGE_HW, 1024 * wm, 512 * wm, NUM_CLASSES)
    else:
        raise ValueError("unknown model kind %r" % kind)
    for i in range(3):                                        # model.py:67-87
        add((fc[i], fc[i + 1]), "fc_w", "fc%d" % (i + 1))
        add((fc[i + 1],), "bias", "fc%d" % (i + 1))
    return ModelSpec(kind, wm, ch, fc, tuple(vs))


REFERENCE_SPEC = make_spec("cnn", 1)
assert REFERENCE_SPEC.num_params == 2_656_010
