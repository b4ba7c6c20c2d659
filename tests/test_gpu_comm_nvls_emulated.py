"""K-AR (NVLS) kernel in its emulated form: every rank on ONE GPU (blockIdx.y = rank) and the multimem
instructions spelled out as peer loads / stores — checks ownership, tails, barriers and the C entry point's
argument checks.  The multicast instructions themselves need >= 2 GPUs behind NVSwitch (tests/mgpu_worker.py
with LDNN_NVLS=1)."""
import ctypes as C

import pytest
import torch

from test_gpu_comm_emulated import Emu, _data

pytestmark = pytest.mark.gpu


def _exchange(L, emu, grads, params=None, mode=0, world=1, weight=None):
    e = L.Exchange()
    gt = emu.table(grads)
    pt = emu.table(params) if params is not None else None
    for i in range(L.MAX_RANKS):
        e.grads[i] = gt[i]
        if pt is not None:
            e.params[i] = pt[i]
    e.n, e.mode = grads[0].numel(), mode
    for i in range(world):
        e.weight[i] = (1.0 / world) if weight is None else weight[i]
    return e


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_nvls_form_reduce_matches_mean(world):
    from ldnn_b200.ops import _lib as L
    n = 4 * 50_021
    emu = Emu(world, n)
    x = _data(world, n, seed=10 + world)
    want = torch.stack(x).double().mean(0)
    e = _exchange(L, emu, x, world=world)
    c = emu.comm()
    L.check(L.lib().ldnn_allreduce_nvls(C.byref(c), C.byref(e), None, None, None))
    emu.ok()
    for r in range(world):
        assert torch.allclose(x[r].double(), want, atol=1e-5), r
        assert torch.equal(x[r], x[0]), r


@pytest.mark.parametrize("world", [2, 8])
@pytest.mark.parametrize("kind", ["adam", "sgd"])
def test_nvls_form_update_tail_equals_two_shot(world, kind):
    """Same inputs through the two-shot kernel and the NVLS-form kernel: identical ownership and optimizer
    arithmetic, so parameters / moments / bf16 shadow agree to summation-order noise."""
    from ldnn_b200.engine import OptimConfig
    from ldnn_b200.ops import _lib as L
    n = 4 * 30_011
    opt = OptimConfig(kind=kind, lr=1e-2)
    p0 = torch.randn(n, generator=torch.Generator().manual_seed(1)).cuda()
    state = {}
    for name in ("two_shot", "nvls"):
        emu = Emu(world, n)
        params = [p0.clone() for _ in range(world)]
        m, v = torch.zeros(n).cuda(), torch.zeros(n).cuda()
        shadow = torch.zeros(n, dtype=torch.bfloat16).cuda()
        for step in range(1, 4):
            grads = _data(world, n, seed=step)
            e = _exchange(L, emu, grads, params, L.X_UPDATE, world)
            e.m, e.v, e.shadow = m.data_ptr(), v.data_ptr(), shadow.data_ptr()
            e.opt = L.make_optim(opt, step, opt.lr)
            c = emu.comm()
            if name == "two_shot":
                L.check(L.lib().ldnn_allreduce_fused(C.byref(c), C.byref(e), None))
            else:
                L.check(L.lib().ldnn_allreduce_nvls(C.byref(c), C.byref(e), None, None, None))
            emu.ok()
        state[name] = (params, m, v, shadow)
    (pa, ma, va, sa), (pb, mb, vb, sb) = state["two_shot"], state["nvls"]
    for r in range(world):
        assert torch.equal(pb[r], pb[0]), r
        diff = (pa[r] - pb[r]).abs()
        assert float((diff > 5e-6 + 1e-5 * pa[r].abs()).float().mean()) < 1e-4, r
    assert torch.allclose(ma, mb, atol=1e-6, rtol=1e-5)
    assert torch.allclose(va, vb, atol=1e-8, rtol=1e-4)


def test_nvls_entry_rejects_what_the_switch_cannot_do():
    from ldnn_b200.ops import _lib as L
    world, n = 4, 4 * 1024
    emu = Emu(world, n)
    x = _data(world, n)
    lib = L.lib()
    e = _exchange(L, emu, x, world=world, weight=[0.4, 0.3, 0.2, 0.1])       # shard weights: not uniform
    c = emu.comm()
    assert lib.ldnn_allreduce_nvls(C.byref(c), C.byref(e), None, None, None) != 0
    assert b"uniform" in lib.ldnn_last_error()
    e = _exchange(L, emu, x, world=world, mode=L.X_SELFMIX)
    assert lib.ldnn_allreduce_nvls(C.byref(c), C.byref(e), None, None, None) != 0
    e = _exchange(L, emu, x, world=world)
    c.emulate = 0                                                            # real mode without a multicast mapping
    assert lib.ldnn_allreduce_nvls(C.byref(c), C.byref(e), None, None, None) != 0
    assert b"multicast" in lib.ldnn_last_error()
