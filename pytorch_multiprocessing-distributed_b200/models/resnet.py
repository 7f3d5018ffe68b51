"""CIFAR-style ResNet family, spec-driven.

Parity target: ref:model/resnet.py:15-125.  Same topology, same parameter shapes and the same
``state_dict`` key names (``conv1.weight``, ``bn1.*``, ``layer{1..4}.{i}.conv{1,2,3}.weight``,
``...bn{1,2,3}.*``, ``...shortcut.0.weight``, ``...shortcut.1.*``, ``linear.{weight,bias}``), so a
checkpoint written by the reference loads here and vice versa (after the ``module.`` prefix the
reference's DDP wrapper adds, see utils/checkpoint.py).

What is different:
  * One ``ResNetSpec`` describes the network; the PyTorch module below and the native step
    engine (parallel/engine.py -> csrc/engine.cu) are both generated from it, so the two can
    never disagree on topology.
  * ``depth 18`` keeps the reference's non-standard ``[1, 1, 1, 1]`` BasicBlock layout
    (ref:model/resnet.py:108-109; 4,903,242 parameters) -- *not* torchvision's ``[2, 2, 2, 2]``.
  * The head averages over whatever spatial extent is left instead of a fixed 4x4 window
    (ref:model/resnet.py:102); identical at 32x32 and it makes 224x224 inputs legal
    (BASELINE.json config 4).
  * The torch forward here is the CPU / oracle path.  On a B200 the trainer runs the native
    engine, which reads and writes the very same parameter tensors (they are views of its slabs).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Iterator, List, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

_LAYOUTS: Dict[int, Tuple[str, Tuple[int, int, int, int]]] = {
    18: ('basic', (1, 1, 1, 1)),
    34: ('basic', (3, 4, 6, 3)),
    50: ('bottleneck', (3, 4, 6, 3)),
    101: ('bottleneck', (3, 4, 23, 3)),
    152: ('bottleneck', (3, 8, 36, 3)),
}
_STAGE_PLANES = (64, 128, 256, 512)
_STAGE_STRIDES = (1, 2, 2, 2)


@dataclass(frozen=True)
class ConvSpec:
    name: str          # state_dict prefix of the conv, e.g. ``layer2.0.conv1``
    bn: str            # state_dict prefix of the BatchNorm that follows it
    cin: int
    cout: int
    ksize: int
    stride: int

    @property
    def pad(self) -> int:
        return self.ksize // 2


@dataclass(frozen=True)
class BlockSpec:
    prefix: str
    kind: str                       # 'basic' | 'bottleneck'
    convs: Tuple[ConvSpec, ...]     # main path, in execution order
    shortcut: ConvSpec | None       # 1x1 projection (conv + BN) or None for identity


@dataclass(frozen=True)
class ResNetSpec:
    depth: int
    kind: str
    blocks_per_stage: Tuple[int, int, int, int]
    num_classes: int = 10
    in_channels: int = 3

    @property
    def expansion(self) -> int:
        return 1 if self.kind == 'basic' else 4

    @property
    def stem(self) -> ConvSpec:
        return ConvSpec('conv1', 'bn1', self.in_channels, 64, 3, 1)

    @property
    def feature_dim(self) -> int:
        return 512 * self.expansion

    def blocks(self) -> List[BlockSpec]:
        out: List[BlockSpec] = []
        cin = 64
        for si, (planes, stride0, n) in enumerate(zip(_STAGE_PLANES, _STAGE_STRIDES, self.blocks_per_stage)):
            for bi in range(n):
                stride = stride0 if bi == 0 else 1
                p = f'layer{si + 1}.{bi}'
                cout = planes * self.expansion
                if self.kind == 'basic':
                    convs = (ConvSpec(f'{p}.conv1', f'{p}.bn1', cin, planes, 3, stride),
                             ConvSpec(f'{p}.conv2', f'{p}.bn2', planes, planes, 3, 1))
                else:
                    convs = (ConvSpec(f'{p}.conv1', f'{p}.bn1', cin, planes, 1, 1),
                             ConvSpec(f'{p}.conv2', f'{p}.bn2', planes, planes, 3, stride),
                             ConvSpec(f'{p}.conv3', f'{p}.bn3', planes, cout, 1, 1))
                sc = None
                if stride != 1 or cin != cout:
                    sc = ConvSpec(f'{p}.shortcut.0', f'{p}.shortcut.1', cin, cout, 1, stride)
                out.append(BlockSpec(p, self.kind, convs, sc))
                cin = cout
        return out

    def all_convs(self) -> List[ConvSpec]:
        """Every conv in ``named_parameters()`` order (stem, then per block main path, shortcut)."""
        res = [self.stem]
        for b in self.blocks():
            res.extend(b.convs)
            if b.shortcut is not None:
                res.append(b.shortcut)
        return res

    def num_parameters(self) -> int:
        n = 0
        for c in self.all_convs():
            n += c.cout * c.cin * c.ksize * c.ksize + 2 * c.cout
        return n + self.feature_dim * self.num_classes + self.num_classes

    def train_flops_per_image(self, hw: int = 32) -> float:
        """2*MACs for fprop + dgrad + wgrad (the stem has no dgrad); SURVEY.md 2.6."""
        total = 0.0
        cur = hw
        sizes: Dict[str, int] = {'conv1': hw}      # spatial size entering each conv
        for b in self.blocks():
            hin = cur
            for c in b.convs:
                sizes[c.name] = hin
                hin = (hin + 2 * c.pad - c.ksize) // c.stride + 1
            if b.shortcut is not None:
                sizes[b.shortcut.name] = cur
            cur = hin
        for c in self.all_convs():
            ho = (sizes[c.name] + 2 * c.pad - c.ksize) // c.stride + 1
            f = 2.0 * ho * ho * c.cout * c.cin * c.ksize * c.ksize
            total += f * (2 if c.name == 'conv1' else 3)
        total += 3 * 2.0 * self.feature_dim * self.num_classes
        return total


def resnet_spec(depth: int = 18, num_classes: int = 10) -> ResNetSpec:
    if depth not in _LAYOUTS:
        raise ValueError(f'unsupported ResNet depth {depth}; choose one of {sorted(_LAYOUTS)}')
    kind, blocks = _LAYOUTS[depth]
    return ResNetSpec(depth, kind, blocks, num_classes)


# ----------------------------------------------------------------------------- torch modules
def _conv(c: ConvSpec) -> nn.Conv2d:
    return nn.Conv2d(c.cin, c.cout, c.ksize, stride=c.stride, padding=c.pad, bias=False)


class ResidualBlock(nn.Module):
    """BasicBlock (2 convs) or Bottleneck (3 convs) built from a BlockSpec; module attribute
    names follow ref:model/resnet.py:18-32 / 46-61 so the keys match."""

    def __init__(self, spec: BlockSpec) -> None:
        super().__init__()
        self.n_convs = len(spec.convs)
        for i, c in enumerate(spec.convs, start=1):
            setattr(self, f'conv{i}', _conv(c))
            setattr(self, f'bn{i}', nn.BatchNorm2d(c.cout))
        self.shortcut = nn.Sequential()
        if spec.shortcut is not None:
            self.shortcut = nn.Sequential(_conv(spec.shortcut), nn.BatchNorm2d(spec.shortcut.cout))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        y = x
        for i in range(1, self.n_convs + 1):
            y = getattr(self, f'bn{i}')(getattr(self, f'conv{i}')(y))
            if i < self.n_convs:
                y = F.relu(y)
        return F.relu(y + self.shortcut(x))


class ResNet(nn.Module):
    def __init__(self, spec: ResNetSpec) -> None:
        super().__init__()
        self.spec = spec
        self.conv1 = _conv(spec.stem)
        self.bn1 = nn.BatchNorm2d(64)
        blocks = spec.blocks()
        pos = 0
        for si, n in enumerate(spec.blocks_per_stage):
            setattr(self, f'layer{si + 1}', nn.Sequential(*[ResidualBlock(b) for b in blocks[pos:pos + n]]))
            pos += n
        self.linear = nn.Linear(spec.feature_dim, spec.num_classes)

    def features(self, x: torch.Tensor) -> torch.Tensor:
        y = F.relu(self.bn1(self.conv1(x)))
        for si in range(4):
            y = getattr(self, f'layer{si + 1}')(y)
        return y.mean(dim=(2, 3))      # == avg_pool2d(out, 4) + flatten at 32x32 input

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.linear(self.features(x))


def build_resnet(depth: int = 18, num_classes: int = 10) -> ResNet:
    return ResNet(resnet_spec(depth, num_classes))


def ResNet18(num_classes: int = 10) -> ResNet:
    """Reference 'ResNet18': BasicBlock [1,1,1,1] (ref:model/resnet.py:108-109)."""
    return build_resnet(18, num_classes)


def ResNet34(num_classes: int = 10) -> ResNet:
    return build_resnet(34, num_classes)


def ResNet50(num_classes: int = 10) -> ResNet:
    return build_resnet(50, num_classes)


def ResNet101(num_classes: int = 10) -> ResNet:
    return build_resnet(101, num_classes)


def ResNet152(num_classes: int = 10) -> ResNet:
    return build_resnet(152, num_classes)


MODEL_ALIASES = {
    'res': 18, 'res18': 18, 'resnet18': 18, 'res34': 34, 'resnet34': 34, 'res50': 50, 'resnet50': 50,
    'res101': 101, 'resnet101': 101, 'res152': 152, 'resnet152': 152,
}


def model_from_name(name: str, num_classes: int = 10) -> ResNet:
    """``--model`` values: the reference accepts only ``res`` (ref:main.py:39-40) and dies with
    UnboundLocalError otherwise; here every depth has a name and anything else is
    a clear error."""
    key = name.lower()
    if key not in MODEL_ALIASES:
        raise ValueError(f"unknown --model '{name}'; choose one of {sorted(MODEL_ALIASES)}")
    return build_resnet(MODEL_ALIASES[key], num_classes)
