"""Execution engines: one interface, two realisations.

``TorchEngine``  — torch ops + autograd on a :class:`SpecNet`; CPU/gloo path and oracle.
``NativeEngine`` — (``models/native.py``) static-graph executor over the hand-written
                   sm_100a kernels, CUDA-graph captured.

Both keep **all learnable state in flat fp32 slabs** (`FlatState`): parameters,
gradients and optimizer moments are single contiguous buffers, with every model
tensor a view at a fixed offset.  That is what lets the exchange be one fused
kernel over a bucket instead of the reference's 65 per-tensor collectives plus 65
elementwise kernels (``Balanced All-Reduce/communication.py:21-31``), and lets the
optimizer run in that kernel's tail (``trainer.py:210`` → K-OPT).
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F

from .models.spec import NetSpec, get_spec
from .models.torch_net import SpecNet, xavier_init_

ALIGN = 64  # elements; every tensor starts on a 256-byte (fp32) / 128-byte (bf16) boundary


@dataclass
class Slot:
    key: str
    shape: Tuple[int, ...]
    offset: int
    numel: int


@dataclass
class OptimConfig:
    kind: str = "adam"         # adam | sgd
    lr: float = 1e-3
    beta1: float = 0.9
    beta2: float = 0.999
    eps: float = 1e-8
    momentum: float = 0.9
    weight_decay: float = 0.0
    lr_step: int = 25
    lr_gamma: float = 0.1


def build_layout(shapes: List[Tuple[str, Tuple[int, ...]]]) -> Tuple[List[Slot], int]:
    slots, off = [], 0
    for key, shp in shapes:
        n = 1
        for s in shp:
            n *= s
        slots.append(Slot(key, tuple(shp), off, n))
        off += (n + ALIGN - 1) // ALIGN * ALIGN
    return slots, off


class FlatState:
    """Flat fp32 parameter / gradient / moment slabs (+ step counters)."""

    def __init__(self, slots: List[Slot], total: int, device, opt: OptimConfig, allocator=None):
        self.slots, self.total, self.opt = slots, total, opt
        alloc = allocator or (lambda n: torch.zeros(n, dtype=torch.float32, device=device))
        self.params = alloc(total)
        self.grads = alloc(total)
        self.m = alloc(total)          # adam exp_avg | sgd momentum buffer
        self.v = alloc(total) if opt.kind == "adam" else None
        self.step = 0                  # optimizer steps taken
        self.epoch = 0                 # scheduler steps taken (StepLR is per *local epoch*, trainer.py:218)
        self.by_key: Dict[str, Slot] = {s.key: s for s in slots}

    def view(self, buf: torch.Tensor, key: str) -> torch.Tensor:
        s = self.by_key[key]
        return buf[s.offset:s.offset + s.numel].view(s.shape)

    @property
    def lr(self) -> float:
        return self.opt.lr * (self.opt.lr_gamma ** (self.epoch // self.opt.lr_step))

    def scheduler_step(self) -> None:
        self.epoch += 1

    # -- reference optimizer math on the flat slabs (CPU path + oracle for K-OPT) --
    @torch.no_grad()
    def optimizer_step(self, grads: Optional[torch.Tensor] = None) -> None:
        g = self.grads if grads is None else grads
        o = self.opt
        self.step += 1
        if o.weight_decay:
            g = g.add(self.params, alpha=o.weight_decay)
        if o.kind == "adam":
            self.m.mul_(o.beta1).add_(g, alpha=1 - o.beta1)
            self.v.mul_(o.beta2).addcmul_(g, g, value=1 - o.beta2)
            bc1 = 1 - o.beta1 ** self.step
            bc2 = 1 - o.beta2 ** self.step
            denom = (self.v.sqrt() / math.sqrt(bc2)).add_(o.eps)
            self.params.addcdiv_(self.m, denom, value=-self.lr / bc1)
        else:
            if self.step == 1:
                self.m.copy_(g)
            else:
                self.m.mul_(o.momentum).add_(g)
            self.params.add_(self.m, alpha=-self.lr)


class Engine:
    """Interface the trainer, evaluator, exchanger and checkpointing rely on."""
    spec: NetSpec
    flat: FlatState
    device: torch.device

    def train_step(self, x, y, update: bool = True):
        """fwd + bwd (+ local optimizer step).  Returns device scalars (loss, n_correct)."""
        raise NotImplementedError

    def eval_step(self, x, y):
        """Returns (loss, n_correct, predictions)."""
        raise NotImplementedError

    def probe_step(self, x, y):
        self.train_step(x, y, update=False)

    def state_dict(self) -> Dict[str, torch.Tensor]:
        raise NotImplementedError

    def load_state_dict(self, sd) -> None:
        raise NotImplementedError

    def params_changed(self) -> None:
        """Called after someone wrote ``flat.params`` behind the engine's back."""

    def scheduler_step(self) -> None:
        self.flat.scheduler_step()

    # optimizer moments travel per key in the reference tensor shape (OIHW for convs),
    # so checkpoints move between engines whose flat layouts differ.
    def export_tensor(self, buf: torch.Tensor, key: str) -> torch.Tensor:
        return self.flat.view(buf, key).detach().cpu().clone()

    def import_tensor(self, buf: torch.Tensor, key: str, t: torch.Tensor) -> None:
        self.flat.view(buf, key).copy_(t.to(buf.device))

    def optimizer_state_dict(self) -> dict:
        f = self.flat
        out = {"kind": f.opt.kind, "step": f.step, "epoch": f.epoch, "m": {}, "v": {}}
        for s in f.slots:
            out["m"][s.key] = self.export_tensor(f.m, s.key)
            if f.v is not None:
                out["v"][s.key] = self.export_tensor(f.v, s.key)
        return out

    def load_optimizer_state_dict(self, sd: dict) -> None:
        f = self.flat
        f.step, f.epoch = int(sd["step"]), int(sd["epoch"])
        if sd.get("kind") != f.opt.kind:
            return
        for s in f.slots:
            if s.key in sd["m"]:
                self.import_tensor(f.m, s.key, sd["m"][s.key])
            if f.v is not None and s.key in sd.get("v", {}):
                self.import_tensor(f.v, s.key, sd["v"][s.key])


def canonical_device(device) -> torch.device:
    d = torch.device(device)
    if d.type == "cuda" and d.index is None:
        d = torch.device("cuda", torch.cuda.current_device())
    return d


class TorchEngine(Engine):
    def __init__(self, spec: NetSpec, device, opt: OptimConfig, init: bool = True):
        self.spec, self.device = spec, canonical_device(device)
        self.model = SpecNet(spec)
        if init:
            xavier_init_(self.model)
        self.model.to(self.device)
        slots, total = build_layout([(k, tuple(p.shape)) for k, p in self.model.named_parameters()])
        self._symm = {}
        self.flat = FlatState(slots, total, self.device, opt, allocator=self._symm_allocator())
        # re-seat every parameter (and its grad) as a view of the flat slabs
        with torch.no_grad():
            for k, p in self.model.named_parameters():
                v = self.flat.view(self.flat.params, k)
                v.copy_(p.data)
                p.data = v
                p.grad = self.flat.view(self.flat.grads, k)

    def _symm_allocator(self):
        """On CUDA the flat slabs live in IPC-shareable memory so the fused P2P exchange can
        map them into the peers; on CPU they are ordinary tensors."""
        if self.device.type != "cuda":
            return None
        from .ops import _lib as L
        if not L.available():
            return None
        from .ops import native_ops as K

        def alloc(n):
            name = ["params", "grads", "m", "v"][len(self._symm)]
            if name in ("params", "grads") and K.nvls_requested():
                t, handle, ptr = K.symm_alloc_multicast(n * 4, self.device)      # multicast-capable slab (NVLS all-reduce); IPC slab if unavailable
            else:
                t, handle, ptr = K.symm_alloc(n * 4, self.device)
            self._symm[name] = (t, handle, ptr)
            return t.view(torch.float32)
        return alloc

    def symm_tensor(self, name: str) -> torch.Tensor:
        return self._symm[name][0]

    def symm_handle(self, name: str):
        if name not in self._symm:
            raise RuntimeError("flat slabs are not in symmetric memory (CPU engine or library not built)")
        return self._symm[name][1], self._symm[name][2], self.flat.total * 4

    def train_step(self, x, y, update: bool = True):
        self.model.train()
        self.flat.grads.zero_()
        logits = self.model(x)
        loss = F.cross_entropy(logits, y)
        loss.backward()
        if update:
            self.flat.optimizer_step()
        with torch.no_grad():
            correct = (logits.argmax(1) == y).sum()
        return loss.detach(), correct

    @torch.no_grad()
    def eval_step(self, x, y):
        self.model.eval()
        logits = self.model(x)
        loss = F.cross_entropy(logits, y)
        pred = logits.argmax(1)
        return loss, (pred == y).sum(), pred

    def state_dict(self):
        return {k: v.detach().clone() for k, v in self.model.state_dict().items()}

    def load_state_dict(self, sd):
        with torch.no_grad():
            own = self.model.state_dict()
            for k, v in sd.items():
                own[k].copy_(v.to(own[k].device))


def make_engine(cfg, rt) -> Engine:
    opt = OptimConfig(kind=cfg.optimizer, lr=cfg.lr, momentum=cfg.momentum, weight_decay=cfg.weight_decay,
                      lr_step=cfg.lr_step, lr_gamma=cfg.lr_gamma)
    kw = {"num_classes": cfg.num_classes}
    if cfg.model == "enhanced_cnn" and os.environ.get("LDNN_TINY_WIDTHS"):
        kw["widths"] = (8, 16, 32, 64, 128)      # test-only: same topology, 1/64 of the FLOPs
    spec = get_spec(cfg.model, **kw)
    kind = cfg.engine
    dtype = getattr(cfg, "dtype", "tf32")
    if kind == "auto":
        # the sm_100a kernels compute in tf32 (fp32 storage, TF32 tensor-core MMA: the reference's numerics) or
        # bf16 (bf16 storage, fp32 master weights); --dtype fp32 selects the torch engine (same flat slabs)
        kind = "native" if (rt.is_cuda and dtype in ("tf32", "bf16")) else "torch"
    if kind == "native":
        if not rt.is_cuda:
            raise RuntimeError("--engine native needs a CUDA device (sm_100a)")
        if dtype not in ("tf32", "bf16"):
            raise RuntimeError("--engine native computes in tf32 or bf16 (fp32 accumulation, fp32 master weights); "
                               "use --engine torch (or auto) with --dtype fp32")
        from .models.native import NativeEngine
        # rings are latency-bound per launch: fewer, larger buckets for the overlapped per-step exchange
        bucket_mb = 64 if (cfg.sync == "step" and cfg.topology != "allreduce") else 32
        return NativeEngine(spec, rt.device, opt, batch_size=cfg.batch_size, world_size=rt.world_size,
                            rank=rt.rank, bucket_mb=bucket_mb, dtype=dtype)
    return TorchEngine(spec, rt.device, opt)
