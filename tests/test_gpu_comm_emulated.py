"""The exchange kernels with every rank emulated on ONE GPU (blockIdx.y = rank, cooperative
launch): exact kernel code, local "peer" buffers.  Oracle = fp64 mixing matrix."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


class Emu:
    def __init__(self, world, n, blocks=4):
        from ldnn_b200.ops import _lib as L
        self.L, self.world, self.n, self.blocks = L, world, n, blocks
        self.sig = [torch.zeros(L.SIGNAL_WORDS, dtype=torch.int32, device="cuda") for _ in range(world)]
        self.err = torch.zeros(1, dtype=torch.int32, device="cuda")
        self.epoch = 0

    def comm(self):
        L = self.L
        self.epoch += 1
        c = L.Comm()
        c.rank, c.world, c.emulate, c.blocks, c.epoch = 0, self.world, 1, self.blocks, self.epoch
        c.timeout_cycles = 4_000_000_000
        for i, s in enumerate(self.sig):
            c.signal[i] = s.data_ptr()
        c.error_flag = self.err.data_ptr()
        return c

    def table(self, tensors):
        arr = (C.c_void_p * self.L.MAX_RANKS)()
        for i, t in enumerate(tensors):
            arr[i] = t.data_ptr()
        return arr

    def ok(self):
        torch.cuda.synchronize()
        assert int(self.err.item()) == 0, "watchdog fired"


def _data(world, n, seed=0):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn(n, generator=g).cuda() for _ in range(world)]


@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("topology", ["allreduce", "ring", "double_ring"])
def test_reduce_matches_oracle(world, topology):
    from ldnn_b200.ops import _lib as L
    from ldnn_b200.parallel.exchange import ExchangeSpec, mixing_matrix, reduce_form
    n = 4 * 50_021            # deliberately not a multiple of world / chunk sizes
    for agg in ["equal", "shard"]:
        emu = Emu(world, n)
        x = _data(world, n, seed=world)
        X = torch.stack(x).double()
        sp = ExchangeSpec(topology, "allreduce", agg, 0.5, [float(i + 1) for i in range(world)], world)
        e = L.Exchange()
        tab = emu.table(x)
        recv = [torch.zeros(n).cuda() for _ in range(world)]
        rtab = emu.table(recv)
        for i in range(L.MAX_RANKS):
            e.grads[i], e.recv[i] = tab[i], rtab[i]
        e.n, e.mode = n, 0
        _, b = reduce_form(sp, 0)
        for i in range(world):
            e.weight[i] = b[i]
        c = emu.comm()
        if topology == "allreduce":
            L.check(L.lib().ldnn_allreduce_fused(C.byref(c), C.byref(e), None))
        else:
            L.check(L.lib().ldnn_ring_fused(C.byref(c), C.byref(e), 1 if topology == "ring" else 2, None))
        emu.ok()
        want = torch.from_numpy(mixing_matrix(sp)).cuda() @ X
        for r in range(world):
            assert torch.allclose(x[r].double(), want[r], atol=1e-5), (topology, agg, r)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("topology", ["allreduce", "ring", "double_ring"])
@pytest.mark.parametrize("kind", ["adam", "sgd"])
def test_fused_update_tail(world, topology, kind):
    """reduce → scale → optimizer → bf16 cast in one kernel == all-reduce then FlatState.optimizer_step."""
    from ldnn_b200.engine import FlatState, OptimConfig, Slot
    from ldnn_b200.ops import _lib as L
    n = 4 * 30_011
    opt = OptimConfig(kind=kind, lr=1e-2)
    emu = Emu(world, n)
    p0 = torch.randn(n).cuda()
    params = [p0.clone() for _ in range(world)]
    ms = [torch.zeros(n).cuda() for _ in range(world)]
    vs = [torch.zeros(n).cuda() for _ in range(world)]
    shadows = [torch.zeros(n, dtype=torch.bfloat16).cuda() for _ in range(world)]
    ref = FlatState([Slot("w", (n,), 0, n)], n, "cuda", opt)
    ref.params.copy_(p0)
    recv = [torch.zeros(n).cuda() for _ in range(world)]
    for step in range(1, 4):
        grads = _data(world, n, seed=step)
        gmean = torch.stack(grads).double().mean(0).float()
        ref.optimizer_step(gmean)
        # emulate mode shares one m/v/shadow pointer per launch: the kernel indexes only the
        # owned slice of m/v, so a single shared m/v is exactly the sharded optimizer state.
        e = L.Exchange()
        gt, pt, rt_ = emu.table(grads), emu.table(params), emu.table(recv)
        for i in range(L.MAX_RANKS):
            e.grads[i], e.params[i], e.recv[i] = gt[i], pt[i], rt_[i]
        e.m, e.v, e.shadow = ms[0].data_ptr(), vs[0].data_ptr(), 0
        e.n, e.mode = n, L.X_UPDATE
        for i in range(world):
            e.weight[i] = 1.0 / world
        e.opt = L.make_optim(opt, step, ref.lr)
        c = emu.comm()
        if topology == "allreduce":
            L.check(L.lib().ldnn_allreduce_fused(C.byref(c), C.byref(e), None))
        else:
            L.check(L.lib().ldnn_ring_fused(C.byref(c), C.byref(e), 1 if topology == "ring" else 2, None))
        emu.ok()
        for r in range(world):
            # Adam's m/(sqrt(v)+eps) is ill-conditioned where the mean gradient is ~0 and the fp32
            # summation order differs from the fp64 oracle: allow a vanishing fraction of outliers.
            diff = (params[r] - ref.params).abs()
            assert float((diff > 5e-6 + 1e-5 * ref.params.abs()).float().mean()) < 1e-4, (step, r)
            assert float(diff.max()) < 2.5e-2, (step, r)
            assert torch.equal(params[r], params[0]), (step, r)
    assert torch.allclose(ms[0], ref.m, atol=1e-6, rtol=1e-5)


@pytest.mark.parametrize("world", [2, 3, 5, 8])
@pytest.mark.parametrize("topology", ["ring", "double_ring"])
@pytest.mark.parametrize("agg", ["equal", "weighted"])
def test_gossip_matches_reference_math(world, topology, agg):
    from ldnn_b200.ops import _lib as L
    from ldnn_b200.parallel.exchange import ExchangeSpec, gossip_coefficients, mixing_matrix
    n = 4 * 20_003
    sp = ExchangeSpec(topology, "gossip", agg, 0.7, None, world)
    x, recv = _data(world, n, 3), [torch.zeros(n).cuda() for _ in range(world)]
    X = torch.stack(x).double()
    emu = Emu(world, n)
    c_self, nbrs = gossip_coefficients(sp, 0)
    offs = (C.c_int * 2)(*([d for d, _ in nbrs] + [0, 0])[:2])
    coefs = (C.c_float * 2)(*([w for _, w in nbrs] + [0.0, 0.0])[:2])
    c = emu.comm()
    L.check(L.lib().ldnn_gossip_pull(C.byref(c), emu.table(x), emu.table(recv), C.c_longlong(n), C.c_float(c_self),
                                     len(nbrs), offs, coefs, None))
    emu.ok()
    want = torch.from_numpy(mixing_matrix(sp)).cuda() @ X
    for r in range(world):
        assert torch.allclose(x[r].double(), want[r], atol=1e-5), (topology, agg, r)


def test_selfmix_weighted_allreduce():
    from ldnn_b200.ops import _lib as L
    from ldnn_b200.parallel.exchange import ExchangeSpec, mixing_matrix, reduce_form
    world, n = 4, 4 * 10_007
    sp = ExchangeSpec("allreduce", "allreduce", "weighted", 0.6, None, world)
    x, recv = _data(world, n, 5), [torch.zeros(n).cuda() for _ in range(world)]
    X = torch.stack(x).double()
    emu = Emu(world, n)
    a, b = reduce_form(sp, 0)
    e = L.Exchange()
    gt, rt_ = emu.table(x), emu.table(recv)
    for i in range(L.MAX_RANKS):
        e.grads[i], e.recv[i] = gt[i], rt_[i]
    e.n, e.mode, e.self_coef = n, L.X_SELFMIX, a
    for i in range(world):
        e.weight[i] = b[i]
    c = emu.comm()
    L.check(L.lib().ldnn_allreduce_fused(C.byref(c), C.byref(e), None))
    emu.ok()
    want = torch.from_numpy(mixing_matrix(sp)).cuda() @ X
    for r in range(world):
        assert torch.allclose(x[r].double(), want[r], atol=1e-5)


def test_broadcast_allgather_barrier_and_repeat():
    from ldnn_b200.ops import _lib as L
    world, n = 8, 4 * 4099
    emu = Emu(world, n)
    bufs = _data(world, n, 9)
    src = bufs[3].clone()
    c = emu.comm()
    L.check(L.lib().ldnn_broadcast(C.byref(c), emu.table(bufs), C.c_size_t(n * 4), 3, None))
    emu.ok()
    assert all(torch.equal(b, src) for b in bufs)
    count = 37
    mine = torch.arange(world * count, dtype=torch.float32).cuda()       # emulate: rank r's record at [r*count]
    tables = [torch.zeros(world * count).cuda() for _ in range(world)]
    for _ in range(5):                                                   # epochs keep increasing, flags never reset
        c = emu.comm()
        L.check(L.lib().ldnn_allgather_small(C.byref(c), C.c_void_p(mine.data_ptr()), emu.table(tables), count, None))
        c = emu.comm()
        L.check(L.lib().ldnn_barrier(C.byref(c), None))
    emu.ok()
    assert all(torch.equal(t, mine) for t in tables)


def test_watchdog_reports_a_missing_peer():
    """Failure detection: a rank whose peer never arrives gets an error flag instead of a hung kernel
    (upstream hangs forever in the next collective)."""
    from ldnn_b200.ops import _lib as L
    emu = Emu(2, 16)
    c = emu.comm()
    c.emulate, c.rank, c.blocks = 0, 0, 2         # real (non-emulated) rank 0 of 2; nobody plays rank 1
    c.timeout_cycles = 60_000_000                # ~30 ms
    L.check(L.lib().ldnn_barrier(C.byref(c), None))
    torch.cuda.synchronize()
    assert int(emu.err.item()) == 1


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("variant", ["selfmix", "update_cast", "nvls_update_cast"])
def test_two_shot_tail_reads_only_what_its_own_barrier_covers(world, variant):
    """The barrier before the two-shot kernels' local tail joins CTA b of this rank with CTA b of each peer only.
    One CTA index of every rank is held back (~1 ms) before it pushes: a tail that walked the slab with another
    element-to-CTA mapping than the pushes would read slices the late CTA has not delivered yet (stale self-mix
    sums, stale bf16 shadow).  Results must be unaffected by the delay."""
    from ldnn_b200.engine import OptimConfig
    from ldnn_b200.ops import _lib as L
    n = 4 * 70_003
    blocks = 6
    opt = OptimConfig(kind="adam", lr=1e-2)
    outs = []
    for delay_block in (0, 3):
        emu = Emu(world, n, blocks=blocks)
        grads = _data(world, n, seed=5)
        p0 = torch.randn(n, generator=torch.Generator().manual_seed(9)).cuda()
        params = [p0.clone() for _ in range(world)]
        recv = [torch.full((n,), 7.0).cuda() for _ in range(world)]       # stale staging content
        m, v = torch.zeros(n).cuda(), torch.zeros(n).cuda()
        shadow = torch.full((n,), -3.0, dtype=torch.bfloat16).cuda()      # stale shadow content
        e = L.Exchange()
        gt, pt, rt_ = emu.table(grads), emu.table(params), emu.table(recv)
        for i in range(L.MAX_RANKS):
            e.grads[i], e.params[i], e.recv[i] = gt[i], pt[i], rt_[i]
        e.n = n
        c = emu.comm()
        c.debug_delay_block, c.debug_delay_cycles = delay_block + (1 if delay_block else 0), 2_000_000
        if variant == "selfmix":
            e.mode, e.self_coef = L.X_SELFMIX, 0.7
            for i in range(world):
                e.weight[i] = 0.3 / max(world - 1, 1)
            L.check(L.lib().ldnn_allreduce_fused(C.byref(c), C.byref(e), None))
            emu.ok()
            outs.append(torch.stack(grads).clone())
        else:
            e.mode = L.X_UPDATE
            e.m, e.v, e.shadow = m.data_ptr(), v.data_ptr(), shadow.data_ptr()
            for i in range(world):
                e.weight[i] = 1.0 / world
            e.opt = L.make_optim(opt, 1, 1e-2)
            if variant == "update_cast":
                L.check(L.lib().ldnn_allreduce_fused(C.byref(c), C.byref(e), None))
            else:
                L.check(L.lib().ldnn_allreduce_nvls(C.byref(c), C.byref(e), None, None, None))
            emu.ok()
            # emulate mode: every rank's tail writes the one shared shadow from its own (identical) parameters
            assert torch.equal(shadow, params[0].bfloat16()), (variant, delay_block)
            outs.append(torch.stack(params).clone())
    assert torch.equal(outs[0], outs[1]), variant


class LLPads:
    """Per-rank staging pads of the flag-free rings (3 phases, sentinel-filled) + the host-side call bookkeeping."""

    def __init__(self, emu, phase4):
        self.emu, self.phase4, self.epoch, self.prev_used = emu, phase4, 0, 0
        self.pads = [torch.full((3 * phase4 * 4,), -2 ** 31, dtype=torch.int32, device="cuda") for _ in range(emu.world)]

    def arm(self, e, n, rings):
        L = self.emu.L
        tab = self.emu.table(self.pads)
        for i in range(L.MAX_RANKS):
            e.ll[i] = tab[i]
        e.ll_phase4, e.ll_clear4, e.ll_epoch = self.phase4, self.prev_used, self.epoch % 3
        self.epoch += 1
        self.prev_used = int(L.lib().ldnn_ring_ll_slots(C.c_longlong(n), self.emu.world, rings))


@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("rings", [1, 2])
def test_flag_free_ring_matches_oracle_over_many_calls(world, rings):
    """ldnn_ring_ll: payload-announces-itself rings.  Nine back-to-back calls of different sizes reuse the three
    phases (each call re-arms the phase the previous one used); inputs contain exact zeros and -0.0, the bit
    pattern the pad uses as 'not here yet'."""
    from ldnn_b200.ops import _lib as L
    from ldnn_b200.parallel.exchange import ExchangeSpec, mixing_matrix, reduce_form
    nmax = 4 * 40_009
    emu = Emu(world, nmax, blocks=4)
    pads = LLPads(emu, 2 * (nmax // 4) + 4096)
    sp = ExchangeSpec("ring" if rings == 1 else "double_ring", "allreduce", "shard", 0.5,
                      [float(i + 1) for i in range(world)], world)
    M = torch.from_numpy(mixing_matrix(sp)).cuda()
    for call, n in enumerate([nmax, 4 * 64, 4 * 1000, nmax, 4 * 9, 4 * 33_333, 4 * 8, nmax, 4 * 2_048]):
        x = _data(world, n, seed=100 + call)
        for t in x:
            t[::7] = 0.0
            t[3::11] = -0.0
        X = torch.stack(x).double()
        e = L.Exchange()
        tab = emu.table(x)
        for i in range(L.MAX_RANKS):
            e.grads[i] = tab[i]
        e.n, e.mode = n, 0
        _, b = reduce_form(sp, 0)
        for i in range(world):
            e.weight[i] = b[i]
        pads.arm(e, n, rings)
        c = emu.comm()
        L.check(L.lib().ldnn_ring_ll(C.byref(c), C.byref(e), rings, None))
        emu.ok()
        want = M @ X
        for r in range(world):
            assert torch.allclose(x[r].double(), want[r], atol=1e-5), (call, n, r)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("rings", [1, 2])
def test_flag_free_ring_fused_update_tail(world, rings):
    """reduce -> scale -> Adam -> bf16 cast inside ldnn_ring_ll == all-reduce then FlatState.optimizer_step."""
    from ldnn_b200.engine import FlatState, OptimConfig, Slot
    from ldnn_b200.ops import _lib as L
    n = 4 * 30_011
    opt = OptimConfig(kind="adam", lr=1e-2)
    emu = Emu(world, n)
    pads = LLPads(emu, 2 * (n // 4) + 4096)
    p0 = torch.randn(n).cuda()
    params = [p0.clone() for _ in range(world)]
    m, v = torch.zeros(n).cuda(), torch.zeros(n).cuda()
    ref = FlatState([Slot("w", (n,), 0, n)], n, "cuda", opt)
    ref.params.copy_(p0)
    for step in range(1, 5):
        grads = _data(world, n, seed=step)
        ref.optimizer_step(torch.stack(grads).double().mean(0).float())
        e = L.Exchange()
        gt, pt = emu.table(grads), emu.table(params)
        for i in range(L.MAX_RANKS):
            e.grads[i], e.params[i] = gt[i], pt[i]
        e.m, e.v, e.shadow = m.data_ptr(), v.data_ptr(), 0
        e.n, e.mode = n, L.X_UPDATE
        for i in range(world):
            e.weight[i] = 1.0 / world
        e.opt = L.make_optim(opt, step, ref.lr)
        pads.arm(e, n, rings)
        c = emu.comm()
        L.check(L.lib().ldnn_ring_ll(C.byref(c), C.byref(e), rings, None))
        emu.ok()
        for r in range(world):
            diff = (params[r] - ref.params).abs()
            assert float((diff > 5e-6 + 1e-5 * ref.params.abs()).float().mean()) < 1e-4, (step, r)
            assert float(diff.max()) < 2.5e-2, (step, r)
            assert torch.equal(params[r], params[0]), (step, r)
    assert torch.allclose(m, ref.m, atol=1e-6, rtol=1e-5)
