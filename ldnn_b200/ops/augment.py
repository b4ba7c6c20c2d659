"""Batched AutoAugment (CIFAR-10 policy) on whole uint8 batches, on whatever device the batch lives on.

The reference augments one PIL image at a time inside a ``DataLoader`` with no workers
(``Balanced All-Reduce/dataloader.py:14-20,42``): ~2 k images/s per process, 25x slower than one B200
trains this model.  Here the shard stays resident in HBM as uint8 (a CIFAR-10 training shard is ~120 MB),
and a chunk of several batches is augmented at once: images are grouped by (operation, magnitude, sign)
— at most 25 sub-policies x 2 operations x 2 signs groups — and each group goes through the *same pixel
math as torchvision* (``torchvision.transforms.autoaugment._apply_op`` on a ``[B, C, H, W]`` tensor), so the
result for a given set of random choices is identical to the per-image transform (tested).  Random choices
follow ``AutoAugment.get_params``: sub-policy uniform, each operation applied when ``u <= p``, sign fair.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch


def batched_equalize(x: torch.Tensor) -> torch.Tensor:
    """Histogram equalisation of every (image, channel) plane of a uint8 ``[B, C, H, W]`` tensor at once.

    torchvision's tensor ``equalize`` loops over images and channels in Python with a data-dependent branch per
    plane (a host synchronisation each on CUDA).  Same integer arithmetic here, vectorised over the B*C planes:
    ``step = floor((pixels - count of the brightest occupied bin) / 255)``,
    ``lut[v] = clamp(floor((cumhist[v-1] + floor(step/2)) / step), 0, 255)``, identity where ``step == 0``."""
    assert x.dtype == torch.uint8 and x.dim() == 4
    B, Cc, H, W = x.shape
    planes = B * Cc
    v = x.reshape(planes, H * W).to(torch.int64)
    offs = torch.arange(planes, device=x.device, dtype=torch.int64).unsqueeze(1) * 256
    hist = torch.bincount((v + offs).reshape(-1), minlength=planes * 256).view(planes, 256)
    occupied = hist != 0
    last = 255 - occupied.flip(1).to(torch.int64).argmax(dim=1)                    # brightest occupied bin
    step = (hist.sum(dim=1) - hist.gather(1, last.unsqueeze(1)).squeeze(1)) // 255
    safe = step.clamp(min=1).unsqueeze(1)
    lut = (hist.cumsum(dim=1) + (step // 2).unsqueeze(1)) // safe
    lut = torch.nn.functional.pad(lut, [1, 0])[:, :-1].clamp(0, 255)
    out = lut.gather(1, v)
    out = torch.where((step == 0).unsqueeze(1), v, out)
    return out.to(torch.uint8).view(B, Cc, H, W)


class BatchedAutoAugment:
    batched = True            # ArrayDataset.batch() hands the whole uint8 batch over in one call

    def __init__(self, policy: str = "cifar10", fill: Optional[List[float]] = None):
        from torchvision.transforms import autoaugment as A
        from torchvision.transforms.functional import InterpolationMode
        pol = {"cifar10": A.AutoAugmentPolicy.CIFAR10, "imagenet": A.AutoAugmentPolicy.IMAGENET,
               "svhn": A.AutoAugmentPolicy.SVHN}[policy]
        self._ref = A.AutoAugment(pol)                     # source of the policy table and magnitude bins
        self._apply_op = A._apply_op
        self.policies = self._ref.policies
        self.interpolation = InterpolationMode.NEAREST
        self.fill = fill
        self._space_cache: Dict[Tuple[int, int], dict] = {}

    # ------------------------------------------------------------------ random choices
    def sample(self, n: int, generator: Optional[torch.Generator] = None, device="cpu"):
        """(sub-policy id [n], uniform draws [n,2], signs [n,2]) — the same three draws as
        ``AutoAugment.get_params``, one set per image."""
        g = generator
        gd = "cpu" if g is None else g.device
        pid = torch.randint(len(self.policies), (n,), generator=g, device=gd)
        u = torch.rand((n, 2), generator=g, device=gd)
        sign = torch.randint(2, (n, 2), generator=g, device=gd)
        return pid.to(device), u.to(device), sign.to(device)

    def _space(self, hw: Tuple[int, int]):
        if hw not in self._space_cache:
            self._space_cache[hw] = self._ref._augmentation_space(10, hw)
        return self._space_cache[hw]

    # ------------------------------------------------------------------ apply
    @torch.no_grad()
    def apply(self, x: torch.Tensor, pid: torch.Tensor, u: torch.Tensor, sign: torch.Tensor) -> torch.Tensor:
        """x: uint8 [N, C, H, W].  Returns a new uint8 tensor; choices as returned by :meth:`sample`."""
        assert x.dtype == torch.uint8 and x.dim() == 4
        out = x.clone()
        # group membership is decided on the host (a few thousand integers): the device never has to answer
        # "is this group empty?", so a CUDA batch is augmented without a single data-dependent synchronisation
        pid, u, sign = pid.cpu(), u.cpu(), sign.cpu()
        space = self._space((x.shape[2], x.shape[3]))
        fill = self.fill
        if isinstance(fill, (int, float)):
            fill = [float(fill)] * x.shape[1]
        for k in (0, 1):                                     # first operation of every sub-policy, then the second
            groups: Dict[Tuple[str, Optional[int]], List[int]] = {}
            for t, sub in enumerate(self.policies):
                name, _, mag_id = sub[k]
                groups.setdefault((name, mag_id), []).append(t)
            p_of = torch.tensor([sub[k][1] for sub in self.policies], dtype=u.dtype, device=u.device)
            active = u[:, k] <= p_of[pid]
            for (name, mag_id), subs in groups.items():
                in_group = active & torch.isin(pid, torch.tensor(subs, device=pid.device))
                if not bool(in_group.any()):
                    continue
                mags, signed = space[name]
                mag = float(mags[mag_id].item()) if mag_id is not None else 0.0
                variants = [(in_group, mag)]
                if signed:
                    variants = [(in_group & (sign[:, k] == 1), mag), (in_group & (sign[:, k] == 0), -mag)]
                for mask, m in variants:
                    idx = mask.nonzero(as_tuple=True)[0]
                    if idx.numel() == 0:
                        continue
                    idx = idx.to(x.device)
                    if name == "Equalize":
                        out[idx] = batched_equalize(out[idx])      # torchvision's version loops per image and channel
                    else:
                        out[idx] = self._apply_op(out[idx], name, m, interpolation=self.interpolation, fill=fill)
        return out

    def __call__(self, x: torch.Tensor, generator: Optional[torch.Generator] = None) -> torch.Tensor:
        pid, u, sign = self.sample(x.shape[0], generator, "cpu")
        return self.apply(x, pid, u, sign)
