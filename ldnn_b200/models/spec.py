"""Network description shared by every execution engine.

The reference hard-codes its network as an ``nn.Module`` class
(``Balanced All-Reduce/model.py:52-111``).  Here a network is *data*: an ordered
list of fused "units" over named activation slots.  The same list drives

* the torch module used on CPU / as the numerical oracle (``models/torch_net.py``),
* the B200 static-graph executor (``models/native.py``), which lowers each unit
  to the hand-written sm_100a kernels and captures the whole step in a CUDA graph,
* the ``state_dict`` key layout (checkpoint compatibility).

A unit is ``conv -> batchnorm [-> + residual] [-> relu]`` — exactly the pattern the
kernels fuse — plus a few structural units (max-pool, the pool+fc+loss head).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional, Tuple


@dataclass(frozen=True)
class ConvBN:
    """conv(k×k, stride, pad, no bias) → BN → (+residual) → (ReLU)."""
    conv_key: str                 # state_dict prefix of the conv ("layer1.0.conv1")
    bn_key: str                   # state_dict prefix of the BN   ("layer1.0.bn1")
    cin: int
    cout: int
    k: int
    stride: int
    pad: int
    src: str                      # input activation slot
    dst: str                      # output activation slot
    relu: bool = True
    residual: Optional[str] = None  # slot added after BN, before ReLU


@dataclass(frozen=True)
class MaxPool:
    k: int
    stride: int
    pad: int
    src: str
    dst: str


@dataclass(frozen=True)
class Head:
    """global-average-pool → linear → (softmax cross-entropy)."""
    fc_key: str
    cin: int
    num_classes: int
    src: str


@dataclass
class NetSpec:
    name: str
    in_channels: int
    in_hw: int
    units: List[object] = field(default_factory=list)

    # ---- derived views -------------------------------------------------
    def conv_units(self) -> List[ConvBN]:
        return [u for u in self.units if isinstance(u, ConvBN)]

    def head(self) -> Head:
        return [u for u in self.units if isinstance(u, Head)][0]

    def param_shapes(self) -> List[Tuple[str, Tuple[int, ...]]]:
        """(key, shape) of every learnable tensor, in ``module.parameters()`` order."""
        out: List[Tuple[str, Tuple[int, ...]]] = []
        for key, kind, shape in self._ordered_entries():
            if kind == "param":
                out.append((key, shape))
        return out

    def state_shapes(self):
        """(key, kind, shape) for every ``state_dict`` entry in torch's order."""
        return list(self._ordered_entries())

    def num_params(self) -> int:
        n = 0
        for _, shp in self.param_shapes():
            c = 1
            for s in shp:
                c *= s
            n += c
        return n

    def slot_shapes(self) -> dict:
        """name -> (H, W, C) of every activation slot, by walking the units (shape inference on the IR)."""
        shp = {"x": (self.in_hw, self.in_hw, self.in_channels)}
        for u in self.units:
            if isinstance(u, Head):
                continue
            H, W, Cc = shp[u.src]
            Ho, Wo = (H + 2 * u.pad - u.k) // u.stride + 1, (W + 2 * u.pad - u.k) // u.stride + 1
            shp[u.dst] = (Ho, Wo, u.cout if isinstance(u, ConvBN) else Cc)
        return shp

    def forward_flops(self, batch: int = 1) -> int:
        """Multiply-add FLOPs (2 per MAC) of the convolutions and the classifier for one forward pass; a training
        step is ~3x (fprop + dgrad + wgrad).  Used for the roofline report."""
        shp = self.slot_shapes()
        total = 0
        for u in self.units:
            if isinstance(u, ConvBN):
                Ho, Wo, _ = shp[u.dst]
                total += 2 * Ho * Wo * u.cout * u.cin * u.k * u.k
            elif isinstance(u, Head):
                total += 2 * u.cin * u.num_classes
        return int(total) * int(batch)

    def _ordered_entries(self):
        # torch orders state_dict by module registration; registration follows the
        # order in which prefixes first appear in ``units`` — except that within a
        # residual block the shortcut is registered after conv2/bn2.  ``order_hint``
        # sorts by (block prefix, rank inside the block).
        ents = []
        for u in self.units:
            if isinstance(u, ConvBN):
                ents.append((u.conv_key, "conv", u))
                ents.append((u.bn_key, "bn", u))
            elif isinstance(u, Head):
                ents.append((u.fc_key, "fc", u))
        ents = _registration_order(ents)
        for key, kind, u in ents:
            if kind == "conv":
                yield (key + ".weight", "param", (u.cout, u.cin, u.k, u.k))
            elif kind == "bn":
                yield (key + ".weight", "param", (u.cout,))
                yield (key + ".bias", "param", (u.cout,))
                yield (key + ".running_mean", "buffer", (u.cout,))
                yield (key + ".running_var", "buffer", (u.cout,))
                yield (key + ".num_batches_tracked", "counter", ())
            else:
                yield (key + ".weight", "param", (u.num_classes, u.cin))
                yield (key + ".bias", "param", (u.num_classes,))


def _registration_order(ents):
    """Move each block's shortcut/downsample entries after its main-path entries."""
    def block_of(key: str) -> str:
        parts = key.split(".")
        return ".".join(parts[:2]) if parts[0].startswith("layer") else key

    out, i = [], 0
    while i < len(ents):
        blk = block_of(ents[i][0])
        j = i
        while j < len(ents) and block_of(ents[j][0]) == blk:
            j += 1
        group = ents[i:j]
        main = [e for e in group if ".shortcut." not in e[0] and ".downsample." not in e[0]]
        side = [e for e in group if e not in main]
        out.extend(main + side)
        i = j
    return out


# ---------------------------------------------------------------------------
# The repo's network: reference ``EnhancedCNNModel`` (Balanced All-Reduce/model.py:74-111)
# ---------------------------------------------------------------------------
def enhanced_cnn_spec(num_classes: int = 10, widths=(64, 128, 256, 512, 1024),
                      in_hw: int = 32) -> NetSpec:
    """3×3 stem, four stages of two residual blocks (first block of a stage
    strides by 2 and has a 1×1 projection shortcut), global pool, linear head.
    44,595,786 parameters in 65 tensors; 128 ``state_dict`` keys."""
    s = NetSpec("enhanced_cnn", 3, in_hw)
    s.units.append(ConvBN("prep.0", "prep.1", 3, widths[0], 3, 1, 1, "x", "a0", relu=True))
    cur, cin, t = "a0", widths[0], 1
    for li, cout in enumerate(widths[1:], start=1):
        for bi in range(2):
            stride = 2 if bi == 0 else 1
            p = f"layer{li}.{bi}"
            mid, out = f"a{t}", f"a{t + 1}"
            s.units.append(ConvBN(f"{p}.conv1", f"{p}.bn1", cin, cout, 3, stride, 1, cur, mid, relu=True))
            if stride != 1 or cin != cout:
                sc = f"s{t}"
                s.units.append(ConvBN(f"{p}.shortcut.0", f"{p}.shortcut.1", cin, cout, 1, stride, 0,
                                      cur, sc, relu=False))
                res = sc
            else:
                res = cur
            s.units.append(ConvBN(f"{p}.conv2", f"{p}.bn2", cout, cout, 3, 1, 1, mid, out,
                                  relu=True, residual=res))
            cur, cin, t = out, cout, t + 2
    s.units.append(Head("fc", cin, num_classes, cur))
    return s


# ---------------------------------------------------------------------------
# Scale-up config (BASELINE.json config 4): ResNet-50 @ 224×224
# ---------------------------------------------------------------------------
def resnet50_spec(num_classes: int = 1000, in_hw: int = 224) -> NetSpec:
    """Bottleneck ResNet-50 with torchvision-compatible ``state_dict`` keys
    (conv1/bn1, layerL.B.conv{1,2,3}/bn{1,2,3}, layerL.0.downsample.{0,1}, fc).
    Not present upstream; added as the large-gradient scale-up model."""
    s = NetSpec("resnet50", 3, in_hw)
    s.units.append(ConvBN("conv1", "bn1", 3, 64, 7, 2, 3, "x", "c1", relu=True))
    s.units.append(MaxPool(3, 2, 1, "c1", "p1"))
    cur, cin, t = "p1", 64, 0
    for li, (planes, blocks, stride) in enumerate([(64, 3, 1), (128, 4, 2), (256, 6, 2), (512, 3, 2)], start=1):
        for bi in range(blocks):
            st = stride if bi == 0 else 1
            p = f"layer{li}.{bi}"
            m1, m2, out = f"b{t}a", f"b{t}b", f"b{t}o"
            cout = planes * 4
            s.units.append(ConvBN(f"{p}.conv1", f"{p}.bn1", cin, planes, 1, 1, 0, cur, m1))
            s.units.append(ConvBN(f"{p}.conv2", f"{p}.bn2", planes, planes, 3, st, 1, m1, m2))
            if bi == 0:
                sc = f"b{t}s"
                s.units.append(ConvBN(f"{p}.downsample.0", f"{p}.downsample.1", cin, cout, 1, st, 0,
                                      cur, sc, relu=False))
                res = sc
            else:
                res = cur
            s.units.append(ConvBN(f"{p}.conv3", f"{p}.bn3", planes, cout, 1, 1, 0, m2, out,
                                  relu=True, residual=res))
            cur, cin, t = out, cout, t + 1
    s.units.append(Head("fc", cin, num_classes, cur))
    return s


SPECS = {"enhanced_cnn": enhanced_cnn_spec, "resnet50": resnet50_spec}


def get_spec(name: str, **kw) -> NetSpec:
    return SPECS[name](**kw)
