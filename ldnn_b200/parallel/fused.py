"""FusedExchanger — the product path of the exchange layer.

Every topology's exchange is one hand-written sm_100a kernel (``csrc/comm.cu``) that reads
and writes the peers' flat slabs directly over NVLink/NVSwitch and runs the per-rank weight,
the optimizer update and the bf16 cast in the same pass:

=============  ===================  =========================================================
topology       semantics            kernel
=============  ===================  =========================================================
allreduce      (always global)      ``ldnn_allreduce_fused`` two-shot: pull-reduce the owned
                                    slice, fused tail, push to every peer
ring           allreduce            ``ldnn_ring_fused(num_rings=1)`` RS+AG around one ring
double_ring    allreduce            ``ldnn_ring_fused(num_rings=2)`` two counter-rotating rings,
                                    each on half of the slab
ring / dring   allreduce, <= 64 MB  ``ldnn_ring_ll``: the same RS+AG schedule, flag-free (payload
                                    announces itself in a sentinel-initialised staging pad)
ring / dring   gossip (reference)   ``ldnn_gossip_pull`` 1-/2-hop neighbour mix, out of place
=============  ===================  =========================================================

Replaces ``average_*`` (BalAR/communication.py:4-31), ``ring_all_reduce_*``
(BalR/communication.py:5-62), ``double_ring_all_reduce*`` (BalDR/communication.py:5-77) and the
per-parameter driver loops (BalR/trainer.py:147-168).  torch.distributed is used only to
exchange the 64-byte CUDA-IPC handles at start-up (control plane).

Self-weighted averaging (``--aggregation_type weighted``) has a rank-dependent result, which a
reduce-scatter cannot carry around a ring; it runs on the two-shot kernel's SELFMIX form for
every topology when ``semantics == "allreduce"``.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import _lib as L
from ..ops import native_ops as K
from .exchange import ExchangeSpec, gossip_coefficients, reduce_form

SM_CLOCK_KHZ = 1_965_000       # fallback when the device does not report a clock (B200 max SM clock)


def _sm_clock_khz(device) -> int:
    """SM clock the watchdog's cycle budget is derived from (cudaDevAttrClockRate, kHz)."""
    try:
        khz = int(getattr(torch.cuda.get_device_properties(device), "clock_rate", 0))
    except Exception:
        khz = 0
    return khz if khz > 0 else SM_CLOCK_KHZ


class SymmetricBuffer:
    """A local slab + the peers' mappings of the same slab (pointer table indexed by rank)."""

    def __init__(self, name: str, local_ptr: int, handle: Optional[bytes], nbytes: int, rt, tensor=None):
        self.name, self.nbytes = name, nbytes
        self.ptrs: List[int] = [0] * rt.world_size
        self.ptrs[rt.rank] = local_ptr
        self.rank, self.device = rt.rank, rt.device
        self._ipc_opened: List[int] = []    # peer mappings this object opened through CUDA IPC (ours to close)
        self.multicast = 0             # multicast address of the slab (NVLS), 0 = none
        if handle is None and rt.world_size > 1:
            # slab from torch's symmetric-memory allocator (LDNN_NVLS=1): peers + multicast from its rendezvous
            if tensor is None:
                raise L.LdnnError(f"symmetric slab `{name}` has neither an IPC handle nor a symmetric-memory tensor")
            self.ptrs, self.multicast, self._hdl = K.symm_rendezvous(tensor)
            if self.ptrs[rt.rank] != local_ptr:
                raise L.LdnnError(f"symmetric slab `{name}`: rendezvous pointer mismatch")
        elif rt.world_size > 1:
            handles: List[Optional[bytes]] = [None] * rt.world_size
            dist.all_gather_object(handles, handle)
            for r, h in enumerate(handles):
                if r != rt.rank:
                    self.ptrs[r] = K.symm_open(h, nbytes, rt.device)
                    self._ipc_opened.append(self.ptrs[r])

    def close_peers(self) -> None:
        """Drop this object's mappings of the peers' slabs (reference-counted by the driver: another exchanger over
        the same slab keeps its own)."""
        with torch.cuda.device(self.device):
            for p in self._ipc_opened:
                L.check(L.lib().ldnn_symm_close(C.c_void_p(p)))
        self._ipc_opened = []

    def free_local(self) -> None:
        """Free the local slab (only for slabs the exchanger allocated itself; call after every peer unmapped it)."""
        p, self.ptrs[self.rank] = self.ptrs[self.rank], 0
        if p:
            with torch.cuda.device(self.device):
                L.check(L.lib().ldnn_symm_free(C.c_void_p(p)))

    def table(self):
        arr = (C.c_void_p * L.MAX_RANKS)()
        for i, p in enumerate(self.ptrs):
            arr[i] = p
        return arr


class FusedExchanger:
    def __init__(self, spec: ExchangeSpec, rt, engine, comm_blocks: int = 0, watchdog_ms: int = 20000):
        self.spec, self.rt, self.engine = spec, rt, engine
        self.rank, self.world = rt.rank, rt.world_size
        if self.world > L.MAX_RANKS:
            raise L.LdnnError(f"fused exchange supports up to {L.MAX_RANKS} ranks on one NVSwitch domain")
        if not hasattr(engine, "symm_handle"):
            raise L.LdnnError("engine does not keep its flat slabs in symmetric memory")
        # measured on 4 and 8 B200 (tools/gpu_tune_overlap.sh): the two-shot kernel saturates NVLink with 32
        # CTAs and leaves more SMs to the overlapped backward; the rings want 64
        self.blocks = comm_blocks or (32 if spec.topology == "allreduce" else 64)
        # the exchange CTAs spin on their peers and are launched non-cooperatively next to the backward kernels: every
        # one of them must be resident at once on every rank, so the grid never exceeds half of the SMs
        try:
            sms = int(torch.cuda.get_device_properties(rt.device).multi_processor_count)
        except Exception:
            sms = 148
        self.blocks = max(1, min(self.blocks, sms // 2))
        self.timeout_cycles = int(watchdog_ms * _sm_clock_khz(rt.device))
        self.epoch = 0
        self.launches = 0
        dev = rt.device
        self._bufs = {}
        for name in ("params", "grads"):
            h, p, nb = engine.symm_handle(name)
            tens = engine.symm_tensor(name) if (h is None and hasattr(engine, "symm_tensor")) else None
            self._bufs[name] = SymmetricBuffer(name, p, h, nb, rt, tensor=tens)
        # in-switch reduce (multimem) is used when both slabs have a multicast mapping (LDNN_NVLS=1 + NVSwitch)
        # The switch-side reduce moves ~S bytes per GPU and direction instead of 2S(W-1)/W: a win that grows with W
        # (8 GPUs: 827 vs 640 GB/s bus bandwidth at 512 MB) and a loss at W = 2, where both move S and the multimem
        # round trip is slower than peer loads (360 vs 600 GB/s): used from LDNN_NVLS_MIN_WORLD ranks up (default 4);
        # LDNN_NVLS=1 forces it at any size.
        self.nvls_capable = all(self._bufs[k].multicast for k in ("params", "grads"))
        self.nvls = self.nvls_capable and (os.environ.get("LDNN_NVLS") == "1" or
                                           self.world >= int(os.environ.get("LDNN_NVLS_MIN_WORLD", "4")))
        t, h, p = K.symm_alloc(L.SIGNAL_WORDS * 4, dev)
        self._signal_t = t
        self._bufs["signal"] = SymmetricBuffer("signal", p, h, L.SIGNAL_WORDS * 4, rt)
        self._recv = None
        # flag-free ring path (ldnn_ring_ll) for payloads up to LDNN_RING_LL_MAX_MB (default 64 = the ring bucket
        # size of the per-step schedule; 0 disables it).  The pad costs 6x that much device memory, allocated on first use
        self.ll_max_bytes = int(float(os.environ.get("LDNN_RING_LL_MAX_MB", "64")) * (1 << 20))
        self._ll = None
        self._ll_epoch = 0              # ldnn_ring_ll calls issued on the pad so far (identical on every rank)
        self._ll_prev_used = 0          # slots the previous call used (re-armed by the next one)
        self.error_flag = torch.zeros(1, dtype=torch.int32, device=dev)
        self.n = engine.flat.total
        if self.world > 1:
            dist.barrier()

    # ------------------------------------------------------------------ plumbing
    def _recv_buf(self) -> SymmetricBuffer:
        if self._recv is None:
            cap = int(self.engine.flat.total)            # sized for the whole slab, whatever bucket asks first
            t, h, p = K.symm_alloc(cap * 4, self.rt.device)
            self._recv_t = t
            self._recv = SymmetricBuffer("recv", p, h, cap * 4, self.rt)
            if self.world > 1:
                dist.barrier()
        return self._recv

    def _ll_buf(self) -> SymmetricBuffer:
        """Staging pad of the flag-free rings: 3 phases, every word = the sentinel 0x80000000 (-0.0f)."""
        if self._ll is None:
            self._ll_phase4 = 2 * (self.ll_max_bytes // 16) + 4096
            nbytes = 3 * self._ll_phase4 * 16
            t, h, p = K.symm_alloc(nbytes, self.rt.device)
            t.view(torch.int32).fill_(-2 ** 31)
            torch.cuda.synchronize(self.rt.device)
            self._ll_t = t
            self._ll = SymmetricBuffer("ll", p, h, nbytes, self.rt)
            if self.world > 1:
                dist.barrier()
        return self._ll

    def _use_ll(self, e: L.Exchange) -> bool:
        return (self.world > 1 and self.spec.topology != "allreduce" and not (e.mode & L.X_SELFMIX)
                and 0 < e.n * 4 <= self.ll_max_bytes)

    def _comm(self) -> L.Comm:
        self.epoch += 1
        c = L.Comm()
        c.rank, c.world, c.emulate, c.blocks = self.rank, self.world, 0, self.blocks
        c.epoch, c.timeout_cycles = self.epoch, self.timeout_cycles
        sig = self._bufs["signal"].table()
        for i in range(L.MAX_RANKS):
            c.signal[i] = sig[i]
        c.error_flag = self.error_flag.data_ptr()
        return c

    def set_shard_weights(self, w) -> None:
        self.spec.shard_weights = list(w)

    def _exchange_struct(self, data: str, mode: int, opt=None, lo: int = 0, hi: Optional[int] = None) -> L.Exchange:
        """Exchange descriptor for flat elements [lo, hi) (a bucket) — the whole slab by default."""
        hi = self.n if hi is None else hi
        e = L.Exchange()
        tab = self._bufs[data].table()
        ptab = self._bufs["params"].table()
        for i in range(self.world):
            e.grads[i], e.params[i] = tab[i] + 4 * lo, ptab[i] + 4 * lo
        f = self.engine.flat
        e.m = f.m.data_ptr() + 4 * lo if f.m is not None else 0
        e.v = f.v.data_ptr() + 4 * lo if f.v is not None else 0
        sh = getattr(self.engine, "shadow", None)
        e.shadow = 0 if sh is None else sh.data_ptr() + 2 * lo
        e.n = hi - lo
        a, b = reduce_form(self.spec, self.rank)
        for i in range(self.world):
            e.weight[i] = b[i]
        e.self_coef = a
        if a != 0.0:
            mode |= L.X_SELFMIX
        if a != 0.0 or (self.spec.topology != "allreduce" and self.world > 1):
            rtab = self._recv_buf().table()          # staging slab: self-mix sums / ring partial sums
            for i in range(self.world):
                e.recv[i] = rtab[i] + 4 * lo
        e.mode = mode
        if opt is not None:
            e.opt = opt
        # multicast addresses of the same bucket, for the NVLS kernel (python-side attributes, not struct fields)
        e.mc_data = self._bufs[data].multicast + 4 * lo if self._bufs[data].multicast else 0
        e.mc_params = self._bufs["params"].multicast + 4 * lo if self._bufs["params"].multicast else 0
        e.uniform = all(abs(b[i] - b[0]) <= 1e-12 for i in range(self.world))
        return e

    def _launch_reduce(self, e: L.Exchange) -> None:
        c = self._comm()
        lib, st = L.lib(), L.stream_ptr()
        if (self.nvls and self.spec.topology == "allreduce" and not (e.mode & L.X_SELFMIX)
                and getattr(e, "uniform", False) and getattr(e, "mc_data", 0)):
            L.check(lib.ldnn_allreduce_nvls(C.byref(c), C.byref(e), C.c_void_p(e.mc_data), C.c_void_p(e.mc_params), st))
        elif self.spec.topology == "allreduce" or (e.mode & L.X_SELFMIX):
            L.check(lib.ldnn_allreduce_fused(C.byref(c), C.byref(e), st))
        else:
            rings = 1 if self.spec.topology == "ring" else 2
            if self._use_ll(e):
                tab = self._ll_buf().table()
                for i in range(self.world):
                    e.ll[i] = tab[i]
                e.ll_phase4, e.ll_clear4, e.ll_epoch = self._ll_phase4, self._ll_prev_used, self._ll_epoch % 3
                L.check(lib.ldnn_ring_ll(C.byref(c), C.byref(e), rings, st))
                self._ll_epoch += 1
                self._ll_prev_used = int(lib.ldnn_ring_ll_slots(C.c_longlong(e.n), self.world, rings))
            else:
                L.check(lib.ldnn_ring_fused(C.byref(c), C.byref(e), rings, st))
        self.launches += 1

    def _gossip(self, data: str) -> None:
        c_self, nbrs = gossip_coefficients(self.spec, self.rank)
        c = self._comm()
        offs = (C.c_int * 2)(*([d for d, _ in nbrs] + [0, 0])[:2])
        coefs = (C.c_float * 2)(*([w for _, w in nbrs] + [0.0, 0.0])[:2])
        L.check(L.lib().ldnn_gossip_pull(C.byref(c), self._bufs[data].table(), self._recv_buf().table(),
                                         C.c_longlong(self.n), C.c_float(c_self), len(nbrs), offs, coefs,
                                         L.stream_ptr()))
        self.launches += 1

    # ------------------------------------------------------------------ engine-level API
    def exchange_grads(self, engine, apply_update: bool = False) -> None:
        f = engine.flat
        if self.spec.is_gossip():
            self._gossip("grads")
            if apply_update:
                engine.local_update() if hasattr(engine, "local_update") else (f.optimizer_step(), engine.params_changed())
            return
        if apply_update:
            f.step += 1
            e = self._exchange_struct("grads", L.X_UPDATE, L.make_optim(f.opt, f.step, f.lr))
            self._update_buckets = [(0, self.n)]
        else:
            e = self._exchange_struct("grads", 0)
        self._launch_reduce(e)
        if apply_update and getattr(engine, "shadow", None) is None:
            engine.params_changed()

    # ------------------------------------------------------------------ bucketed, overlapped with backward
    def begin_step(self, engine):
        """Start an overlapped step: fixes the optimizer scalars for every bucket of this step."""
        f = engine.flat
        f.step += 1
        self._step_opt = L.make_optim(f.opt, f.step, f.lr)
        self._update_buckets = []
        if not hasattr(self, "_comm_stream"):
            self._comm_stream = torch.cuda.Stream(self.rt.device, priority=-1)

    def exchange_bucket(self, engine, lo: int, hi: int) -> None:
        """Called right after the backward kernels that complete flat-grad range [lo, hi) were
        enqueued on the current stream: the fused exchange+update for that bucket runs on the
        high-priority comm stream while backward continues."""
        main = torch.cuda.current_stream(self.rt.device)
        ev = torch.cuda.Event()
        ev.record(main)
        self._comm_stream.wait_event(ev)
        with torch.cuda.stream(self._comm_stream):
            e = self._exchange_struct("grads", L.X_UPDATE, self._step_opt, lo, hi)
            self._launch_reduce(e)
        self._update_buckets.append((lo, hi))

    def end_step(self, engine) -> None:
        """The next forward needs every bucket's updated parameters."""
        ev = torch.cuda.Event()
        ev.record(self._comm_stream)
        torch.cuda.current_stream(self.rt.device).wait_event(ev)

    def gather_moments(self, engine) -> bool:
        """Collective, between steps.  The fused tails run the optimizer only on the slice a rank owns, so after
        per-step fused updates every rank holds current moments for its own slices only (a sharded optimizer
        state).  Before the state is exported (checkpoint) the slices are put back together on every rank.

        Ownership is not re-derived here: the same kernels are replayed on the same buckets with a probe
        optimizer (SGD, lr 0, momentum 0: ``m <- g``, parameters unchanged) on an all-ones gradient, which
        marks exactly the elements this rank updates — for whichever kernel (two-shot, rings, NVLS) and bucket
        layout the step used.  The marks must tile the slab (checked).  Returns True if a gather ran.
        Uses ``torch.distributed`` for the two reductions (control plane; checkpoint time only)."""
        f = engine.flat
        buckets = list(getattr(self, "_update_buckets", []) or [])
        if self.world == 1 or not buckets or f.m is None:
            return False
        dev = self.rt.device
        if dev.type == "cuda":
            torch.cuda.synchronize(dev)
        probe = torch.zeros_like(f.m)
        f.grads.fill_(1.0)                       # the gradient slab is dead between steps
        opt = L.Optim()
        opt.kind, opt.lr, opt.momentum, opt.weight_decay, opt.grad_scale = 1, 0.0, 0.0, 0.0, 1.0
        opt.beta1 = opt.beta2 = opt.eps = 0.0
        opt.bc1 = opt.bc2 = 1.0
        for lo, hi in buckets:
            e = self._exchange_struct("grads", L.X_UPDATE, opt, lo, hi)
            e.m, e.v, e.shadow = probe.data_ptr() + 4 * lo, 0, 0
            self._launch_reduce(e)
        self.check()
        mine = probe != 0
        cover = mine.to(torch.float32)
        dist.all_reduce(cover)
        f.grads.zero_()
        self._update_buckets = []
        if not bool((cover > 0).any()):
            raise L.LdnnError("moment gather: the probe marked nothing although update buckets were recorded")
        if bool(((cover == float(self.world)) | (cover == 0)).all()):
            return False                         # every rank updated everything (self-weighted form): replicated
        if bool((cover > 1).any()):
            raise L.LdnnError("moment gather: two ranks claim the same optimizer slice")
        untouched = cover == 0                   # padding outside every bucket: identical everywhere, keep
        for t in (f.m, f.v):
            if t is None:
                continue
            t.copy_(torch.where(mine, t, torch.where(untouched, t / self.world, torch.zeros_like(t))))
            dist.all_reduce(t)
        return True

    def exchange_weights(self, engine) -> None:
        if self.spec.is_gossip():
            self._gossip("params")
        else:
            self._launch_reduce(self._exchange_struct("params", L.X_CAST))
        if getattr(engine, "shadow", None) is None or self.spec.is_gossip():
            engine.params_changed()

    def broadcast_state(self, engine, src: int = 0) -> None:
        """C1: flat parameters by one P2P pull; BN buffers (tiny) over the control plane."""
        if self.world == 1:
            return
        c = self._comm()
        L.check(L.lib().ldnn_broadcast(C.byref(c), self._bufs["params"].table(), C.c_size_t(self.n * 4), src,
                                       L.stream_ptr()))
        self.launches += 1
        for name in ("running_mean", "running_var", "nbt"):
            t = getattr(engine, name, None)
            if t is not None:
                dist.broadcast(t, src=src)
        if hasattr(engine, "model"):
            for b in engine.model.buffers():
                dist.broadcast(b, src=src)
        engine.params_changed()
        self.check()

    # ------------------------------------------------------------------ K-MET: small all-gather
    METRIC_CAP = 4096          # floats per rank

    def allgather_floats(self, rec):
        """Every rank's float record lands in slot [rank] of every peer's table (one P2P kernel;
        replaces the reference's per-epoch scalar collectives C3-C5, C11, C13)."""
        import numpy as np
        n = int(len(rec))
        if n > self.METRIC_CAP:
            raise ValueError(f"metric record of {n} floats exceeds the table capacity {self.METRIC_CAP}")
        if not hasattr(self, "_met"):
            t, h, p = K.symm_alloc(self.world * self.METRIC_CAP * 4, self.rt.device)
            self._met_t = t.view(torch.float32)
            self._met = SymmetricBuffer("metrics", p, h, self.world * self.METRIC_CAP * 4, self.rt)
            self._met_mine = torch.zeros(self.METRIC_CAP, dtype=torch.float32, device=self.rt.device)
            if self.world > 1:
                dist.barrier()
        self._met_mine[:n].copy_(torch.as_tensor(np.asarray(rec, dtype=np.float32)), non_blocking=True)
        c = self._comm()
        L.check(L.lib().ldnn_allgather_small(C.byref(c), L.ptr(self._met_mine), self._met.table(), n, L.stream_ptr()))
        self.launches += 1
        torch.cuda.current_stream(self.rt.device).synchronize()
        return self._met_t[: self.world * n].view(self.world, n).cpu().numpy()

    # ------------------------------------------------------------------ host-visible flag words (C6)
    USER_BASE = L.SIGNAL_WORDS - 64

    def flag_views(self):
        """(my slot in every PEER's signal pad, my own pad's slots [world]) as int32 device tensors: the last 64
        words of the pad are never touched by the kernels.  A store into a peer view crosses NVLink."""
        pad = self._bufs["signal"]
        peers = [K.tensor_from_ptr(pad.ptrs[p] + 4 * (self.USER_BASE + self.rank), 4, torch.int32, self.rt.device)
                 for p in range(self.world) if p != self.rank]
        mine = K.tensor_from_ptr(pad.ptrs[self.rank] + 4 * self.USER_BASE, 4 * self.world, torch.int32, self.rt.device)
        return peers, mine

    def barrier(self) -> None:
        c = self._comm()
        L.check(L.lib().ldnn_barrier(C.byref(c), L.stream_ptr()))

    def check(self) -> None:
        """Raise if a device-side watchdog fired (a peer never arrived)."""
        torch.cuda.synchronize(self.rt.device)
        if int(self.error_flag.item()) != 0:
            raise L.LdnnError("exchange watchdog expired: a peer did not reach the flag in time")

    def close(self) -> None:
        """Collective.  Unmaps the peers' slabs and frees what this exchanger allocated (signal pad, staging slabs,
        metric table); the engine's parameter/gradient slabs stay with the engine.  Order: everyone idle -> everyone
        unmaps -> everyone frees (freeing memory a peer still maps is the failure mode)."""
        if getattr(self, "_closed", False):
            return
        self._closed = True
        torch.cuda.synchronize(self.rt.device)
        if self.world > 1:
            dist.barrier()
        own = [b for b in (self._bufs.get("signal"), self._recv, self._ll, getattr(self, "_met", None)) if b is not None]
        try:
            for b in list(self._bufs.values()) + own[1:]:
                b.close_peers()
            if self.world > 1:
                dist.barrier()
            for name in ("_signal_t", "_recv_t", "_ll_t", "_met_t"):      # non-owning views of the slabs
                if hasattr(self, name):
                    setattr(self, name, None)
            for b in own:
                b.free_local()
        except L.LdnnError as err:                                          # teardown must not take a run down
            import warnings
            warnings.warn(f"FusedExchanger.close: {err}")
        self._recv = self._ll = None
        if hasattr(self, "_met"):
            del self._met


# ---------------------------------------------------------------------------------------------------------------
# Fused kernels for ARBITRARY CUDA tensors (the reference-signature communication.py / trainer.py entry points)
# ---------------------------------------------------------------------------------------------------------------
class _StageEngine:
    """Engine facade around one symmetric staging slab (what FusedExchanger needs: flat slabs + IPC handles)."""

    def __init__(self, n: int, device):
        from types import SimpleNamespace
        self._symm, self.shadow = {}, None
        self.flat = SimpleNamespace(total=n, m=None, v=None)
        for name in ("params", "grads"):            # `params` is never exchanged here: a 16-byte stand-in
            t, h, p = K.symm_alloc((n if name == "grads" else 4) * 4, device)
            self._symm[name] = (t, h, p, (n if name == "grads" else 4) * 4)
            setattr(self.flat, name, t.view(torch.float32))

    def symm_handle(self, name):
        return self._symm[name][1], self._symm[name][2], self._symm[name][3]

    def params_changed(self):
        pass


class StagedFusedExchange:
    """``exchange_(tensor, spec)`` on the fused P2P kernels for any fp32 CUDA tensor: the tensor is copied into a
    symmetric staging slab, exchanged there by the topology's kernel (two-shot / rings / gossip pull, exactly the
    kernels the engines use on their own slabs) and copied back.  This is what the variant directories'
    ``communication.py`` functions and the reference-signature ``train_global`` run on a single-node GPU job, so those
    entry points are not an NCCL path.  Collective: every rank must call with the same sizes in the same order (the
    reference's per-parameter loops do).  The slab grows (collectively) to the largest tensor seen."""

    _instances = {}

    @classmethod
    def get(cls, rt):
        key = (str(rt.device), rt.rank, rt.world_size)
        if key not in cls._instances:
            cls._instances[key] = cls(rt)
        return cls._instances[key]

    def __init__(self, rt):
        self.rt, self.cap, self.eng, self.ex = rt, 0, None, {}

    def _ensure(self, n: int):
        if n <= self.cap:
            return
        for ex in self.ex.values():
            ex.close()
        self.ex = {}
        self.cap = max(1 << 20, 1 << (n - 1).bit_length())
        self.eng = _StageEngine(self.cap, self.rt.device)

    def _exchanger(self, spec: ExchangeSpec) -> FusedExchanger:
        if spec.topology not in self.ex:            # CTA budget depends on the topology only
            self.ex[spec.topology] = FusedExchanger(spec, self.rt, self.eng)
        ex = self.ex[spec.topology]
        ex.spec = spec
        return ex

    @torch.no_grad()
    def exchange_(self, t: torch.Tensor, spec: ExchangeSpec) -> torch.Tensor:
        n = t.numel()
        npad = (n + 3) // 4 * 4
        self._ensure(npad)
        slab = self.eng.flat.grads
        slab[:n].copy_(t.reshape(-1))
        if npad > n:
            slab[n:npad].zero_()
        ex = self._exchanger(spec)
        ex.n = npad
        if spec.is_gossip():
            ex._gossip("grads")
        else:
            ex._launch_reduce(ex._exchange_struct("grads", 0))
        t.copy_(slab[:n].view_as(t))
        return t

    @classmethod
    def close_all(cls) -> None:
        """Collective: unmap and free every staging slab / signal pad the reference-signature entry points created
        (call on every rank before the process group is destroyed; a later exchange builds them again)."""
        for inst in list(cls._instances.values()):
            for ex in inst.ex.values():
                ex.close()
            inst.ex, inst.eng, inst.cap = {}, None, 0
        cls._instances.clear()

    @classmethod
    def usable(cls, rt, t: torch.Tensor) -> bool:
        """Single-node NCCL job, fp32 CUDA tensor, at most LDNN_MAX_RANKS ranks, library present."""
        if not (t.is_cuda and t.dtype == torch.float32 and dist.is_available() and dist.is_initialized()):
            return False
        ws = dist.get_world_size()
        if ws < 2 or ws > L.MAX_RANKS or dist.get_backend() != "nccl" or os.environ.get("LDNN_COMPAT_DIST"):
            return False
        if int(os.environ.get("LOCAL_WORLD_SIZE", str(ws))) != ws:
            return False
        return L.available()
