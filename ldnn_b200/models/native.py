"""NativeEngine: static-graph executor of a :class:`NetSpec` on the hand-written sm_100a kernels.

What torch autograd + cuDNN do for the reference (``trainer.py:205-210`` through
``model.py:102-111``) is done here by an explicit plan:

* all buffers are allocated once (activations NHWC, flat fp32 params/grads/moments
  in IPC-shareable memory, weights in KRSC order);
* two precisions (``dtype``): ``"tf32"`` — fp32 activations and the fp32 master weights fed straight to
  ``tcgen05.mma kind::tf32`` (the reference's own numerics: fp32 storage + TF32 convolutions, no shadow
  copy, no cast) — and ``"bf16"`` — bf16 activations and a bf16 compute copy of the weights (``kind::f16``);
* forward = per unit ``conv_fprop → bn_stats → bn_apply(+residual,+relu)``, head =
  ``pool+fc+CE (+dlogits, d_act, fc grads)``; backward walks the units in reverse:
  ``bn_bwd_reduce → bn_bwd_apply → conv_wgrad (→ flat grad slab) → conv_dgrad``;
  gradient fan-in at residual joins is folded into the dgrad / bn_bwd epilogues
  (accumulate flags decided at plan time) — there is no ``zero_grad`` and no add kernel;
* the whole step is captured into one CUDA graph per batch size; the optimizer (or the
  fused exchange+optimizer kernel) is the only other launch of a step.

The flat parameter layout is *internal* (KRSC, stem rows zero-padded to a multiple of 64);
``state_dict()`` / ``load_state_dict()`` translate to the reference's OIHW fp32 layout.
"""
from __future__ import annotations

import math
import os
from typing import Dict, List, Optional, Tuple

import torch

from ..engine import ALIGN, Engine, FlatState, OptimConfig, Slot, canonical_device
from ..ops import _lib as L
from ..ops import native_ops as K
from .spec import ConvBN, Head, MaxPool, NetSpec
from .torch_net import SpecNet, xavier_init_

BN_EPS, BN_MOMENTUM = 1e-5, 0.1


def _pad64(v: int) -> int:
    return (v + 63) // 64 * 64


class _ConvInfo:
    """Static facts about one ConvBN unit."""

    def __init__(self, u: ConvBN, H: int, W: int):
        self.u = u
        self.H, self.W = H, W
        self.P = (H + 2 * u.pad - u.k) // u.stride + 1
        self.Q = (W + 2 * u.pad - u.k) // u.stride + 1
        self.im2col = (u.cin % 64) != 0          # 3-channel stem → patch matrix + plain GEMM
        self.kdim = _pad64(u.k * u.k * u.cin) if self.im2col else u.k * u.k * u.cin
        self.w_numel = u.cout * self.kdim


class NativeEngine(Engine):
    permuted_layout = True

    def __init__(self, spec: NetSpec, device, opt: OptimConfig, batch_size: int = 64, world_size: int = 1,
                 rank: int = 0, init: bool = True, use_graph: bool = True, bucket_mb: Optional[int] = None,
                 dtype: str = "tf32"):
        if dtype not in ("tf32", "bf16"):
            raise ValueError(f"NativeEngine dtype must be 'tf32' or 'bf16', got {dtype!r}")
        self.dtype = dtype
        self.act_dtype = torch.float32 if dtype == "tf32" else torch.bfloat16
        self.dcode = 1 if dtype == "tf32" else 0          # ldnn_conv_desc.dtype
        if not L.available():
            raise L.LdnnError("libldnn_b200.so is not built; run `python -m ldnn_b200.build`")
        self.spec, self.device, self.max_batch = spec, canonical_device(device), batch_size
        self.use_graph = use_graph and not os.environ.get("LDNN_NO_GRAPH")
        # BN statistics: fused into the conv epilogue with fp32 red.add (fast, summation order not
        # fixed) or the ticketed two-level reduction (bit-reproducible, ~2x more launches).
        self.deterministic = bool(os.environ.get("LDNN_DETERMINISTIC"))
        self.bucket_elems = int(os.environ.get("LDNN_BUCKET_MB", "32")) * (1 << 20) // 4
        if bucket_mb and "LDNN_BUCKET_MB" not in os.environ:
            self.bucket_elems = bucket_mb * (1 << 20) // 4
        # graph topology: wgrad runs on a side stream concurrently with the dgrad -> bn-backward chain
        # (both only read dy_raw), and the local optimizer runs bucket by bucket on a third stream
        # while backward continues.  Off in deterministic mode (shared split-K scratch).
        self.concurrent = not self.deterministic and not os.environ.get("LDNN_SERIAL")
        self._opt_stream = None
        self.autotune = not os.environ.get("LDNN_NO_AUTOTUNE")
        self.graph_kernel_launches = 0    # kernels launched through CUDA-graph replays (the library's
                                          # own counter only sees eager launches)
        torch.cuda.set_device(self.device)
        # ---- geometry ------------------------------------------------------------
        self.info: Dict[str, _ConvInfo] = {}
        self.slot_shape: Dict[str, Tuple[int, int, int]] = {"x": (spec.in_hw, spec.in_hw, spec.in_channels)}
        for u in spec.units:
            if isinstance(u, ConvBN):
                H, W, _ = self.slot_shape[u.src]
                ci = _ConvInfo(u, H, W)
                self.info[u.conv_key] = ci
                self.slot_shape[u.dst] = (ci.P, ci.Q, u.cout)
            elif isinstance(u, MaxPool):
                H, W, Cc = self.slot_shape[u.src]
                self.slot_shape[u.dst] = ((H + 2 * u.pad - u.k) // u.stride + 1, (W + 2 * u.pad - u.k) // u.stride + 1, Cc)
        # ---- flat layout (internal order == module.parameters() order) -----------
        slots: List[Slot] = []
        off = 0
        for key, shape in spec.param_shapes():
            base = key.rsplit(".", 1)[0]
            n = self.info[base].w_numel if (base in self.info and key.endswith(".weight")) else math.prod(shape)
            slots.append(Slot(key, tuple(shape), off, n))
            off += (n + ALIGN - 1) // ALIGN * ALIGN
        self._symm = {}          # name -> (tensor, ipc handle, raw ptr)
        self.flat = FlatState(slots, off, self.device, opt, allocator=self._alloc_flat)
        # bf16 compute copy of the weights; the tf32 path reads the fp32 master slab directly
        self.shadow = torch.zeros(off, dtype=torch.bfloat16, device=self.device) if dtype == "bf16" else None
        # ---- BN buffers -------------------------------------------------------------
        self.bn_off: Dict[str, int] = {}
        c_tot = 0
        for u in spec.conv_units():
            self.bn_off[u.bn_key] = c_tot
            c_tot += u.cout
        z = lambda n, dt=torch.float32: torch.zeros(n, dtype=dt, device=self.device)
        self.running_mean, self.running_var = z(c_tot), torch.ones(c_tot, dtype=torch.float32, device=self.device)
        self.nbt = z(len(self.bn_off), torch.int64)
        self.save_mean, self.save_invstd, self.scale, self.shift = z(c_tot), z(c_tot), z(c_tot), z(c_tot)
        self.bn_sums = z(4 * c_tot)       # [fwd: per BN [2][C]] [bwd: per BN [2][C]]; zeroed once per step
        self.c_tot = c_tot
        self._plans: Dict[Tuple[int, bool], "_StepPlan"] = {}
        if init:
            ref = SpecNet(spec)
            xavier_init_(ref)
            self.load_state_dict(ref.state_dict())

    # ------------------------------------------------------------------ memory
    def _alloc_flat(self, n: int) -> torch.Tensor:
        name = ["params", "grads", "m", "v"][len(self._symm)]
        if name in ("params", "grads") and K.nvls_requested():
            t, handle, ptr = K.symm_alloc_multicast(n * 4, self.device)     # multicast-capable slab (NVLS all-reduce); IPC slab if unavailable
        else:
            t, handle, ptr = K.symm_alloc(n * 4, self.device)
        self._symm[name] = (t, handle, ptr)
        return t.view(torch.float32)

    def symm_handle(self, name: str):
        return self._symm[name][1], self._symm[name][2], self.flat.total * 4

    def symm_tensor(self, name: str) -> torch.Tensor:
        return self._symm[name][0]

    # ------------------------------------------------------------------ state_dict <-> internal layout
    def _to_internal(self, key: str, t: torch.Tensor) -> torch.Tensor:
        base = key.rsplit(".", 1)[0]
        if base in self.info and key.endswith(".weight"):
            ci = self.info[base]
            w = t.permute(0, 2, 3, 1).reshape(ci.u.cout, -1)             # OIHW → [Cout][R*S*Cin]
            if ci.kdim != w.shape[1]:
                w = torch.nn.functional.pad(w, (0, ci.kdim - w.shape[1]))
            return w.reshape(-1)
        return t.reshape(-1)

    def _from_internal(self, key: str, flat: torch.Tensor) -> torch.Tensor:
        s = self.flat.by_key[key]
        base = key.rsplit(".", 1)[0]
        if base in self.info and key.endswith(".weight"):
            ci = self.info[base]
            u = ci.u
            w = flat.view(u.cout, ci.kdim)[:, : u.k * u.k * u.cin]
            return w.reshape(u.cout, u.k, u.k, u.cin).permute(0, 3, 1, 2).contiguous()
        return flat.view(s.shape).clone()

    def export_tensor(self, buf, key):
        s = self.flat.by_key[key]
        return self._from_internal(key, buf[s.offset:s.offset + s.numel]).detach().cpu()

    def import_tensor(self, buf, key, t):
        s = self.flat.by_key[key]
        buf[s.offset:s.offset + s.numel].copy_(self._to_internal(key, t.to(self.device, torch.float32)))

    def state_dict(self):
        out = {}
        i = 0
        for key, kind, shape in self.spec.state_shapes():
            if kind == "param":
                s = self.flat.by_key[key]
                out[key] = self._from_internal(key, self.flat.params[s.offset:s.offset + s.numel])
            else:
                bn = key.rsplit(".", 1)[0]
                o, c = self.bn_off[bn], shape[0] if shape else 0
                if key.endswith("running_mean"):
                    out[key] = self.running_mean[o:o + c].clone()
                elif key.endswith("running_var"):
                    out[key] = self.running_var[o:o + c].clone()
                else:
                    out[key] = self.nbt[list(self.bn_off).index(bn)].clone()
        return out

    @torch.no_grad()
    def load_state_dict(self, sd):
        bn_keys = list(self.bn_off)
        for key, kind, shape in self.spec.state_shapes():
            if key not in sd:
                continue
            t = sd[key]
            if kind == "param":
                self.import_tensor(self.flat.params, key, t)
            else:
                bn = key.rsplit(".", 1)[0]
                o = self.bn_off[bn]
                if key.endswith("running_mean"):
                    self.running_mean[o:o + t.numel()].copy_(t.to(self.device))
                elif key.endswith("running_var"):
                    self.running_var[o:o + t.numel()].copy_(t.to(self.device))
                else:
                    self.nbt[bn_keys.index(bn)] = int(t)
        self.params_changed()

    def params_changed(self) -> None:
        if self.shadow is not None:
            K.cast_bf16(self.flat.params, self.shadow, self.flat.total)

    # ------------------------------------------------------------------ steps
    def _plan(self, n: int, training: bool) -> "_StepPlan":
        key = (n, training)
        if key not in self._plans:
            self._plans[key] = _StepPlan(self, n, training)
        return self._plans[key]

    def train_step(self, x, y, update: bool = True):
        p = self._plan(int(x.shape[0]), True)
        if update and self.concurrent:
            self._train_step_overlapped_update(p, x, y)
        else:
            p.run(x, y)
            if update:
                self.local_update()
        return p.loss[0], p.correct[0]

    def _train_step_overlapped_update(self, p, x, y) -> None:
        """Local schedule: the optimizer of each finished gradient bucket runs on its own stream while
        the rest of backward is still executing (Adam is HBM-bound, backward is latency-bound)."""
        f = self.flat
        f.step += 1
        o = L.make_optim(f.opt, f.step, f.lr)
        if self._opt_stream is None:
            self._opt_stream = torch.cuda.Stream(self.device)
        main = torch.cuda.current_stream(self.device)
        opt_stream = self._opt_stream

        def on_bucket(lo, hi):
            ev = torch.cuda.Event()
            ev.record(main)
            opt_stream.wait_event(ev)
            with torch.cuda.stream(opt_stream):
                K.optim_step(f.params[lo:hi], f.grads[lo:hi], f.m[lo:hi], None if f.v is None else f.v[lo:hi],
                             None if self.shadow is None else self.shadow[lo:hi], hi - lo, o)

        p.run_bucketed(x, y, on_bucket)
        ev = torch.cuda.Event()
        ev.record(opt_stream)
        main.wait_event(ev)

    def train_step_sync(self, x, y, exchanger):
        """Product schedule (--sync step): forward/backward with the fused exchange+update of each
        gradient bucket overlapped with the rest of backward (BASELINE north star)."""
        if getattr(exchanger, "world", 0) == 1:
            return self.train_step(x, y, update=True)      # no peers: the local bucketed optimizer is the same update
        p = self._plan(int(x.shape[0]), True)
        exchanger.begin_step(self)
        p.run_bucketed(x, y, lambda lo, hi: exchanger.exchange_bucket(self, lo, hi))
        exchanger.end_step(self)
        return p.loss[0], p.correct[0]

    def local_update(self) -> None:
        f = self.flat
        f.step += 1
        o = L.make_optim(f.opt, f.step, f.lr)
        K.optim_step(f.params, f.grads, f.m, f.v, self.shadow, f.total, o)     # shadow None (tf32): no cast

    @torch.no_grad()
    def eval_step(self, x, y):
        p = self._plan(int(x.shape[0]), False)
        p.run(x, y)
        return p.loss[0], p.correct[0], p.pred[: x.shape[0]]

    def launches_per_step(self, n: Optional[int] = None) -> int:
        return self._plan(n or self.max_batch, True).num_launches


class _StepPlan:
    """Buffers + launch list for one (batch size, train/eval) pair; CUDA-graph captured."""

    def __init__(self, eng: NativeEngine, n: int, training: bool):
        self.eng, self.n, self.training = eng, n, training
        dev, spec = eng.device, eng.spec
        bf = lambda *shape: torch.zeros(*shape, dtype=eng.act_dtype, device=dev)     # activation storage type
        f32 = lambda *shape: torch.zeros(*shape, dtype=torch.float32, device=dev)
        H0, W0, C0 = eng.slot_shape["x"]
        self.x_in = f32(n, C0, H0, W0)
        self.labels = torch.zeros(n, dtype=torch.int64, device=dev)
        self.act: Dict[str, torch.Tensor] = {}
        self.dact: Dict[str, torch.Tensor] = {}
        for name, (H, W, Cc) in eng.slot_shape.items():
            if name == "x":
                continue
            self.act[name] = bf(n, H, W, Cc)
            if training:
                self.dact[name] = bf(n, H, W, Cc)
        self.yraw: Dict[str, torch.Tensor] = {}
        self.col: Dict[str, torch.Tensor] = {}
        max_raw = 0
        for u in spec.conv_units():
            ci = eng.info[u.conv_key]
            self.yraw[u.conv_key] = bf(n, ci.P, ci.Q, u.cout)
            max_raw = max(max_raw, n * ci.P * ci.Q * u.cout)
            if ci.im2col:
                self.col[u.conv_key] = bf(n, ci.P, ci.Q, ci.kdim)
        self.dy_raw = bf(max_raw) if training else None
        self.dy_raw_unit = {}
        if training and eng.concurrent:       # wgrad(k) on the side stream outlives unit k: private dy_raw per unit
            for u in spec.conv_units():
                ci = eng.info[u.conv_key]
                self.dy_raw_unit[u.conv_key] = bf(n * ci.P * ci.Q * u.cout)
        self.side = torch.cuda.Stream(dev) if (training and eng.concurrent) else None
        hd = spec.head()
        self.pooled, self.logits, self.dlogits = f32(n, hd.cin), f32(n, hd.num_classes), f32(n, hd.num_classes)
        self.loss_n, self.correct_n = f32(n), f32(n)
        self.pred = torch.zeros(n, dtype=torch.int64, device=dev)
        self.loss, self.correct = f32(1), f32(1)
        self.ops: List = []           # closures, in launch order
        self.num_launches = 0
        self._build()
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.seg_graphs: Optional[List[torch.cuda.CUDAGraph]] = None
        self._warm = False

    # -- helpers -----------------------------------------------------------------
    def _chunks(self, M: int, C: int = 64) -> int:
        """Row chunks of the BN reductions: (C/64 channel slabs) x chunks CTAs, ~2 waves at most."""
        return max(1, min(M // 128, max(1, 296 // max(1, C // 64))))

    def _add(self, fn, launches: int = 1, side: bool = False):
        self.ops.append((fn, side and self.side is not None))
        self.num_launches += launches

    def _build(self):
        eng, n, spec = self.eng, self.n, self.eng.spec
        fl = eng.flat
        P = lambda key: fl.params[fl.by_key[key].offset:fl.by_key[key].offset + fl.by_key[key].numel]
        W = P if eng.shadow is None else (                 # conv weights as the MMA reads them
            lambda key: eng.shadow[fl.by_key[key].offset:fl.by_key[key].offset + fl.by_key[key].numel])
        G = lambda key: fl.grads[fl.by_key[key].offset:fl.by_key[key].offset + fl.by_key[key].numel]
        dev = eng.device
        # ---- first pass: size the shared split-K workspace ----
        descs = []
        for u in spec.conv_units():
            ci = eng.info[u.conv_key]
            geo = (n, ci.P, ci.Q, ci.kdim, u.cout, 1, 1, 0) if ci.im2col else (n, ci.H, ci.W, u.cin, u.cout, u.k, u.stride, u.pad)
            descs.append((K.FPROP, geo))
            if self.training:
                descs.append((K.WGRAD, geo))
                if u.src != "x":
                    descs.append((K.DGRAD, geo))
        ws = K.Workspace(dev)
        atomic = not eng.deterministic
        need = 0
        for m, g in descs:
            need = max(need, K.conv_query(m, *g, atomic_out=(atomic and m == K.WGRAD), dtype=eng.dcode)[0])
            if n != eng.max_batch:                   # this size replays the nominal batch's tuning (see _tuned)
                cfg = K.tuned_conv_config(m, (eng.max_batch,) + tuple(g[1:]), None, False, eng.dcode)
                try:
                    need = max(need, K.conv_query(m, *g, block_n=cfg[0], splits=cfg[1],
                                                  atomic_out=(atomic and m == K.WGRAD), dtype=eng.dcode)[0])
                except L.LdnnError:
                    pass
            if eng.autotune and n == eng.max_batch:  # the tuner may pick any candidate: size for the largest
                n_total = g[4] if m == K.FPROP else g[3]
                for bn, sp in K.conv_candidates(m, g, n_total):
                    try:
                        need = max(need, K.conv_query(m, *g, block_n=bn, splits=sp,
                                                      atomic_out=(atomic and m == K.WGRAD), dtype=eng.dcode)[0])
                    except L.LdnnError:
                        pass
        ws.reserve(min(need, 1 << 30))
        self.ws = ws
        partial_max = 0
        for u in spec.conv_units():
            ci = eng.info[u.conv_key]
            partial_max = max(partial_max, self._chunks(n * ci.P * ci.Q, u.cout) * 2 * u.cout)
        self.partial = torch.zeros(partial_max, dtype=torch.float32, device=dev)
        self.bn_ticket = torch.zeros(4, dtype=torch.int32, device=dev)

        # ---- forward ----
        self.pool_idx = {}
        self.zero_ranges = []          # flat-grad ranges that atomic split-K wgrads accumulate into
        zero_slot = len(self.ops)
        if self.training and not eng.deterministic:
            self._add(lambda: eng.bn_sums.zero_(), 0)          # memset node, not one of our kernels
        for u in spec.units:
            if isinstance(u, ConvBN):
                self._fwd_unit(u, W, P)
            elif isinstance(u, MaxPool):
                H, Wd, Cc = eng.slot_shape[u.src]
                xs, ys = self.act[u.src], self.act[u.dst]
                ix = torch.zeros(ys.numel(), dtype=torch.uint8, device=dev) if self.training else None
                self.pool_idx[u.dst] = ix
                self._add(lambda xs=xs, ys=ys, ix=ix, H=H, Wd=Wd, Cc=Cc, u=u: K.maxpool_fwd(xs, ys, ix, n, H, Wd, Cc, u.k,
                                                                                        u.stride, u.pad))
            elif isinstance(u, Head):
                H, Wd, Cc = eng.slot_shape[u.src]
                a = self.act[u.src]
                da = self.dact[u.src] if self.training else None
                fw, fb = P(u.fc_key + ".weight"), P(u.fc_key + ".bias")
                gw = G(u.fc_key + ".weight") if self.training else None
                gb = G(u.fc_key + ".bias") if self.training else None
                self._add(lambda a=a, da=da, fw=fw, fb=fb, gw=gw, gb=gb, H=H, Wd=Wd, Cc=Cc, u=u: K.head_fwd_bwd(
                    a, n, H * Wd, Cc, u.num_classes, fw, fb, self.labels, self.pooled, self.logits, self.dlogits,
                    self.loss_n, self.correct_n, self.pred, da, gw, gb, self.loss, self.correct, self.training),
                    2 if u.num_classes <= 64 else (5 if self.training else 4))
        if not self.training:
            return
        # ---- backward ----
        # Buckets for the overlapped exchange: flat grads are laid out in forward order and produced in
        # reverse, so finished gradients always form a suffix [lo, total); a bucket closes at a residual
        # block boundary once it holds >= bucket_elems elements.
        self.bucket_marks = []           # (number of ops enqueued so far, lo, hi)
        done_keys = {spec.head().fc_key + ".weight", spec.head().fc_key + ".bias"}
        prev_lo = fl.total

        def close_bucket(force=False):
            nonlocal prev_lo
            lo = fl.total
            for sl in reversed(fl.slots):            # longest finished suffix
                if sl.key not in done_keys:
                    break
                lo = sl.offset
            if lo < prev_lo and (force or prev_lo - lo >= eng.bucket_elems):
                self.bucket_marks.append((len(self.ops), lo, prev_lo))
                prev_lo = lo

        # BN-backward reduce fused into the dgrad that writes the LAST contribution to a slot's gradient:
        # the first consumer of a slot in forward order is the last writer in backward order.
        self.fuse_into = {}              # consumer conv_key -> owner unit (whose BN reduce it absorbs)
        self.fused_owner = set()
        # Measured on B200 (batch 64): the fused variant is ~4 % SLOWER for this net — the dgrad epilogue
        # is on the critical path of a handful of CTAs, while the stand-alone reduce spreads over the
        # whole chip — so it is opt-in (LDNN_BWD_FUSE=1) and the default keeps the separate kernel.
        if not eng.deterministic and os.environ.get("LDNN_BWD_FUSE"):
            owner = {u.dst: u for u in spec.conv_units()}
            first_consumer = {}
            for u in spec.units:
                for slot in ([getattr(u, "src", None)] + ([u.residual] if isinstance(u, ConvBN) and u.residual else [])):
                    if slot is not None and slot not in first_consumer:
                        first_consumer[slot] = u
            for slot, cu in first_consumer.items():
                if slot in owner and isinstance(cu, ConvBN) and cu.src == slot and not eng.info[cu.conv_key].im2col:
                    self.fuse_into[cu.conv_key] = owner[slot]
                    self.fused_owner.add(owner[slot].conv_key)
        written = {spec.head().src}      # slots whose gradient buffer already holds a contribution
        for u in reversed(spec.units):
            if isinstance(u, ConvBN):
                self._bwd_unit(u, W, P, G, written)
                done_keys.update({u.conv_key + ".weight", u.bn_key + ".weight", u.bn_key + ".bias"})
                close_bucket()
            elif isinstance(u, MaxPool):
                H, Wd, Cc = eng.slot_shape[u.src]
                assert u.src not in written, "max-pool input with gradient fan-in is not supported"
                ix, dys, dxs = self.pool_idx[u.dst], self.dact[u.dst], self.dact[u.src]
                self._add(lambda ix=ix, dys=dys, dxs=dxs, H=H, Wd=Wd, Cc=Cc, u=u: K.maxpool_bwd(
                    ix, dys, dxs, n, H, Wd, Cc, u.k, u.stride, u.pad))
                written.add(u.src)
        close_bucket(force=True)
        assert prev_lo == 0, "backward did not produce every gradient"
        # one memset per merged range (ranges of neighbouring layers are fused across small gaps)
        merged = []
        for a, b in sorted(self.zero_ranges):
            if merged and a - merged[-1][1] <= (1 << 16):
                merged[-1][1] = max(merged[-1][1], b)
            else:
                merged.append([a, b])
        for a, b in merged:
            view = fl.grads[a:b]
            self.ops.insert(zero_slot, (view.zero_, False))
            self.bucket_marks = [(k + 1, lo, hi) for k, lo, hi in self.bucket_marks]

    def _tuned(self, mode, geo, make_plan, allow=True):
        """(block_n, splits, variant) for one conv of this plan.  Only the engine's nominal batch size is tuned by
        timing (and served from the shipped tunings): a training run meets a new partial-batch size after almost
        every re-shard, and timing ~70 candidate plans per layer for each of them cost seconds per global epoch
        (the `Balanced Ring` entrypoint fell from 83 k to 35 k samples/s that way in round 1).  Other batch sizes
        reuse the nominal size's choice when it builds for them, else the library heuristics."""
        eng = self.eng
        if self.n == eng.max_batch or os.environ.get("LDNN_TUNE_ALL_BATCHES"):
            return K.tuned_conv_config(mode, geo, make_plan, eng.autotune and allow, eng.dcode)
        nominal = (eng.max_batch,) + tuple(geo[1:])
        cfg = K.tuned_conv_config(mode, nominal, None, False, eng.dcode)
        if cfg != (0, 0, 0):
            try:
                make_plan(*cfg)
                return cfg
            except L.LdnnError:
                pass
        return (0, 0, 0)

    def _fwd_unit(self, u: ConvBN, W, P):
        eng, n = self.eng, self.n
        ci = eng.info[u.conv_key]
        yraw = self.yraw[u.conv_key]
        w = W(u.conv_key + ".weight")
        if ci.im2col:
            col = self.col[u.conv_key]
            H0, W0, C0 = eng.slot_shape["x"]
            self._add(lambda: K.im2col_stem(self.x_in, col, n, C0, H0, W0, u.k, u.stride, u.pad, ci.kdim))
        M, Cc = n * ci.P * ci.Q, u.cout
        o = eng.bn_off[u.bn_key]
        gamma, beta = P(u.bn_key + ".weight"), P(u.bn_key + ".bias")
        sc, sh = eng.scale[o:o + Cc], eng.shift[o:o + Cc]
        rm, rv = eng.running_mean[o:o + Cc], eng.running_var[o:o + Cc]
        sm, si = eng.save_mean[o:o + Cc], eng.save_invstd[o:o + Cc]
        nbt = eng.nbt[list(eng.bn_off).index(u.bn_key):][:1]
        res = self.act[u.residual] if u.residual is not None else None
        out = self.act[u.dst]
        fast = not eng.deterministic
        fsum = eng.bn_sums[2 * o:2 * o + 2 * Cc] if (fast and self.training) else None
        if ci.im2col:
            geo, xin = (n, ci.P, ci.Q, ci.kdim, u.cout, 1, 1, 0), self.col[u.conv_key]
        else:
            geo, xin = (n, ci.H, ci.W, u.cin, u.cout, u.k, u.stride, u.pad), self.act[u.src]
        bn_, sp_, lt_ = self._tuned(
            K.FPROP, geo, lambda b, s_, l_: K.ConvPlan(K.FPROP, *geo, xin, w, yraw, bn_sums=fsum, block_n=b, splits=s_,
                                                       **K.variant_kwargs(l_)))
        plan = K.ConvPlan(K.FPROP, *geo, xin, w, yraw, workspace=self.ws, bn_sums=fsum, block_n=bn_, splits=sp_,
                          **K.variant_kwargs(lt_))
        self._add(plan.run, plan.num_launches)
        if fast:
            self._add(lambda: K.bn_fwd_apply_sums(yraw, fsum, gamma, beta, rm, rv, nbt, sm, si, res, out, M, Cc, u.relu,
                                                  self.training, BN_EPS, BN_MOMENTUM))
            return
        if self.training:
            chunks = self._chunks(M, Cc)
            self._add(lambda: K.bn_fwd_stats(yraw, M, Cc, chunks, self.partial, self.bn_ticket, gamma, beta, rm, rv,
                                             nbt, sm, si, sc, sh, BN_EPS, BN_MOMENTUM))
        else:
            self._add(lambda: K.bn_eval_coeffs(gamma, beta, rm, rv, sc, sh, Cc, BN_EPS))
        self._add(lambda: K.bn_fwd_apply(yraw, sc, sh, res, out, M, Cc, u.relu))

    def _bwd_unit(self, u: ConvBN, W, P, G, written):
        eng, n = self.eng, self.n
        ci = eng.info[u.conv_key]
        M, Cc = n * ci.P * ci.Q, u.cout
        o = eng.bn_off[u.bn_key]
        yraw, out, dout = self.yraw[u.conv_key], self.act[u.dst], self.dact[u.dst]
        assert u.dst in written, f"no gradient reaches {u.dst}"
        sm, si = eng.save_mean[o:o + Cc], eng.save_invstd[o:o + Cc]
        gamma = P(u.bn_key + ".weight")
        dgamma, dbeta = G(u.bn_key + ".weight"), G(u.bn_key + ".bias")
        chunks = self._chunks(M, Cc)
        dyr = self.dy_raw_unit.get(u.conv_key, self.dy_raw[:M * Cc])
        dres, res_acc = None, False
        if u.residual is not None:
            dres = self.dact[u.residual]
            res_acc = u.residual in written
            written.add(u.residual)
        if eng.deterministic:
            self._add(lambda: K.bn_bwd_reduce(dout, out, yraw, sm, si, self.partial, self.bn_ticket[1:], M, Cc,
                                              chunks, u.relu, dgamma, dbeta))
            self._add(lambda: K.bn_bwd_apply(dout, out, yraw, sm, si, gamma, dgamma, dbeta, dyr, dres, M, Cc, u.relu,
                                             res_acc))
        else:
            bsum = eng.bn_sums[2 * eng.c_tot + 2 * o:2 * eng.c_tot + 2 * o + 2 * Cc]
            if u.conv_key not in self.fused_owner:      # else: already reduced in the epilogue of the last dgrad
                self._add(lambda: K.bn_bwd_reduce_sums(dout, out, yraw, sm, si, bsum, M, Cc, chunks, u.relu))
            self._add(lambda: K.bn_bwd_apply_sums(dout, out, yraw, sm, si, gamma, bsum, dyr, dres, dgamma, dbeta, M,
                                                  Cc, u.relu, res_acc))
        w = W(u.conv_key + ".weight")
        gw = G(u.conv_key + ".weight")
        if ci.im2col:
            geo = (n, ci.P, ci.Q, ci.kdim, u.cout, 1, 1, 0)
            xin = self.col[u.conv_key]
        else:
            geo = (n, ci.H, ci.W, u.cin, u.cout, u.k, u.stride, u.pad)
            xin = self.act[u.src]
        atomic = not eng.deterministic
        bn_, sp_, lt_ = self._tuned(
            K.WGRAD, geo, lambda b, s_, l_: K.ConvPlan(K.WGRAD, *geo, xin, None, dyr, dw=gw, atomic_out=atomic,
                                                       block_n=b, splits=s_, **K.variant_kwargs(l_)), allow=atomic)
        wg = K.ConvPlan(K.WGRAD, *geo, xin, None, dyr, dw=gw, workspace=self.ws, atomic_out=atomic, block_n=bn_,
                        splits=sp_, **K.variant_kwargs(lt_))
        if atomic and K.conv_query(K.WGRAD, *geo, block_n=bn_, splits=sp_, dtype=eng.dcode)[0] > 0:   # split-K → red.add into a zeroed slice
            sl = eng.flat.by_key[u.conv_key + ".weight"]
            self.zero_ranges.append((sl.offset, sl.offset + sl.numel))
        self._add(wg.run, wg.num_launches, side=True)
        if u.src == "x":
            return
        dx = self.dact[u.src]
        acc = u.src in written
        # strided 1x1 dgrads leave parity classes untouched: start from zero and accumulate
        full_cover = K.conv_query(K.DGRAD, *geo, accumulate=True, dtype=eng.dcode)[3] == u.stride * u.stride
        if not acc and not full_cover:
            self._add(lambda: dx.zero_(), 0)
            acc = True
        bn_, sp_, lt_ = self._tuned(
            K.DGRAD, geo, lambda b, s_, l_: K.ConvPlan(K.DGRAD, *geo, dx, w, dyr, addend=dx, accumulate=True, block_n=b,
                                                       splits=s_, **K.variant_kwargs(l_)))
        fuse = None
        if u.conv_key in self.fuse_into:
            ow = self.fuse_into[u.conv_key]
            oo = eng.bn_off[ow.bn_key]
            fuse = (self.act[ow.dst], self.yraw[ow.conv_key], eng.save_mean[oo:oo + ow.cout],
                    eng.save_invstd[oo:oo + ow.cout],
                    eng.bn_sums[2 * eng.c_tot + 2 * oo:2 * eng.c_tot + 2 * oo + 2 * ow.cout], ow.relu)
        dg = K.ConvPlan(K.DGRAD, *geo, dx, w, dyr, addend=dx if acc else None, accumulate=acc, workspace=self.ws,
                        block_n=bn_, splits=sp_, bn_bwd_fuse=fuse, **K.variant_kwargs(lt_))
        self._add(dg.run, dg.num_launches)
        written.add(u.src)

    # -- execution ---------------------------------------------------------------
    def _launch_range(self, a: int, b: int):
        """Enqueue ops[a:b]; side-stream ops fork after the preceding main-stream op and are joined
        back at the end of the range (so a range is a self-contained graph segment)."""
        main = torch.cuda.current_stream(self.eng.device)
        forked = False
        for fn, side in self.ops[a:b]:
            if side:
                ev = torch.cuda.Event()
                ev.record(main)
                self.side.wait_event(ev)
                with torch.cuda.stream(self.side):
                    fn()
                forked = True
            else:
                fn()
        if forked:
            ev = torch.cuda.Event()
            ev.record(self.side)
            main.wait_event(ev)

    def _launch_all(self):
        self._launch_range(0, len(self.ops))

    def run(self, x: torch.Tensor, y: torch.Tensor):
        self.x_in.copy_(x, non_blocking=True)
        self.labels.copy_(y, non_blocking=True)
        if not self.eng.use_graph:
            self._launch_all()
            return
        if self.graph is None:
            if not self._warm:
                # one eager pass first: lazy kernel attributes, tickets exercised once
                self._launch_all()
                torch.cuda.current_stream().synchronize()
                self._warm = True
                return
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._launch_all()
            self.graph = g
        self.graph.replay()
        self.eng.graph_kernel_launches += self.num_launches

    def run_bucketed(self, x: torch.Tensor, y: torch.Tensor, on_bucket):
        """Same step, cut into one CUDA graph per bucket: after segment k is enqueued,
        ``on_bucket(lo, hi)`` is called so the exchange / optimizer of that flat-grad range can start on
        another stream while segment k+1 (the rest of backward) runs."""
        self.x_in.copy_(x, non_blocking=True)
        self.labels.copy_(y, non_blocking=True)
        bounds = [0] + [k for k, _, _ in self.bucket_marks]
        if not self._warm:
            self._launch_all()
            torch.cuda.current_stream().synchronize()
            self._warm = True
            for _, lo, hi in self.bucket_marks:
                on_bucket(lo, hi)
            return
        if self.seg_graphs is None and self.eng.use_graph:
            self.seg_graphs = []
            for a, b in zip(bounds[:-1], bounds[1:]):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    self._launch_range(a, b)
                self.seg_graphs.append(g)
        for i, (_, lo, hi) in enumerate(self.bucket_marks):
            if self.eng.use_graph:
                self.seg_graphs[i].replay()
            else:
                self._launch_range(bounds[i], bounds[i + 1])
            on_bucket(lo, hi)
        if self.eng.use_graph:
            self.eng.graph_kernel_launches += self.num_launches
