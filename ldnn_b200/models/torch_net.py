"""Torch-op realisation of a :class:`NetSpec`.

Used (a) as the CPU / gloo execution path (BASELINE config 1), (b) as the fp32
numerical oracle for every sm_100a kernel test, (c) to prove ``state_dict``
layout compatibility with the reference network
(``Balanced All-Reduce/model.py:74-111``: 128 keys, OIHW fp32 weights).

The module tree is generated from the spec's dotted keys, so
``SpecNet(enhanced_cnn_spec()).state_dict().keys()`` is exactly the reference's
key list in the reference's order.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .spec import ConvBN, Head, MaxPool, NetSpec, enhanced_cnn_spec, resnet50_spec


class _Node(nn.Module):
    """Pure container: children are added by dotted path."""


def _attach(root: nn.Module, dotted: str, leaf: nn.Module) -> None:
    parts = dotted.split(".")
    cur = root
    for p in parts[:-1]:
        if p not in cur._modules:
            cur.add_module(p, _Node())
        cur = cur._modules[p]
    cur.add_module(parts[-1], leaf)


def _lookup(root: nn.Module, dotted: str) -> nn.Module:
    cur = root
    for p in dotted.split("."):
        cur = cur._modules[p]
    return cur


class SpecNet(nn.Module):
    def __init__(self, spec: NetSpec):
        super().__init__()
        self.spec = spec
        # Register in torch's state_dict order (shortcut after the main path).
        seen = set()
        by_key = {}
        for u in spec.units:
            if isinstance(u, ConvBN):
                by_key[u.conv_key] = nn.Conv2d(u.cin, u.cout, u.k, u.stride, u.pad, bias=False)
                by_key[u.bn_key] = nn.BatchNorm2d(u.cout)
            elif isinstance(u, Head):
                by_key[u.fc_key] = nn.Linear(u.cin, u.num_classes)
        for key, _, _ in spec.state_shapes():
            mod_key = key.rsplit(".", 1)[0]
            if mod_key not in seen:
                seen.add(mod_key)
                _attach(self, mod_key, by_key[mod_key])

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        slots = {"x": x}
        for u in self.spec.units:
            if isinstance(u, ConvBN):
                y = _lookup(self, u.bn_key)(_lookup(self, u.conv_key)(slots[u.src]))
                if u.residual is not None:
                    y = y + slots[u.residual]
                slots[u.dst] = F.relu(y) if u.relu else y
            elif isinstance(u, MaxPool):
                slots[u.dst] = F.max_pool2d(slots[u.src], u.k, u.stride, u.pad)
            elif isinstance(u, Head):
                pooled = slots[u.src].mean(dim=(2, 3))
                return _lookup(self, u.fc_key)(pooled)
        raise RuntimeError("spec has no head")


def xavier_init_(model: nn.Module) -> None:
    """Reference init: Xavier-uniform conv/linear weights, zero biases
    (``Balanced All-Reduce/main.py:33-37``); BN stays at torch defaults."""
    for m in model.modules():
        if isinstance(m, (nn.Conv2d, nn.Linear)):
            nn.init.xavier_uniform_(m.weight)
            if m.bias is not None:
                nn.init.zeros_(m.bias)


def EnhancedCNNModel(num_classes: int = 10) -> SpecNet:
    """Drop-in for the reference class of the same name."""
    return SpecNet(enhanced_cnn_spec(num_classes))


def ResNet50(num_classes: int = 1000) -> SpecNet:
    return SpecNet(resnet50_spec(num_classes))


class ResBlock(nn.Module):
    """Stand-alone residual block with the reference's constructor signature
    (``Balanced All-Reduce/model.py:52-72``), for users who compose their own nets."""

    def __init__(self, in_channels: int, out_channels: int, stride: int = 1):
        super().__init__()
        self.conv1 = nn.Conv2d(in_channels, out_channels, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(out_channels)
        self.conv2 = nn.Conv2d(out_channels, out_channels, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(out_channels)
        self.shortcut = nn.Sequential()
        if stride != 1 or in_channels != out_channels:
            self.shortcut = nn.Sequential(nn.Conv2d(in_channels, out_channels, 1, stride, bias=False),
                                          nn.BatchNorm2d(out_channels))

    def forward(self, x):
        y = F.relu(self.bn1(self.conv1(x)))
        y = self.bn2(self.conv2(y)) + self.shortcut(x)
        return F.relu(y)
