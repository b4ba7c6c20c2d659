"""One eager training step + emulated (single-GPU, 4 ranks) exchange kernels, for ncu captures of the
non-conv kernel families (BN, head, optimizer, exchange)."""
import ctypes as C, os, sys
os.environ["LDNN_NO_GRAPH"] = "1"; os.environ["LDNN_NO_AUTOTUNE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from ldnn_b200.engine import OptimConfig
from ldnn_b200.models import enhanced_cnn_spec
from ldnn_b200.models.native import NativeEngine
from ldnn_b200.ops import _lib as L
from test_gpu_comm_emulated import Emu, _data
eng = NativeEngine(enhanced_cnn_spec(), "cuda:0", OptimConfig(), batch_size=64)
x = torch.randn(64, 3, 32, 32, device="cuda"); y = torch.randint(0, 10, (64,), device="cuda")
for _ in range(2):
    eng.train_step(x, y)
torch.cuda.synchronize()
world, n = 4, 4 * 1_000_000
emu = Emu(world, n, blocks=16)
for topo in ("allreduce", "ring", "double_ring"):
    grads, params, recv = _data(world, n, 1), _data(world, n, 2), [torch.zeros(n).cuda() for _ in range(world)]
    m, v = torch.zeros(n).cuda(), torch.zeros(n).cuda()
    shadow = torch.zeros(n, dtype=torch.bfloat16).cuda()
    e = L.Exchange()
    gt, pt, rt = emu.table(grads), emu.table(params), emu.table(recv)
    for i in range(L.MAX_RANKS):
        e.grads[i], e.params[i], e.recv[i] = gt[i], pt[i], rt[i]
    e.m, e.v, e.shadow, e.n, e.mode = m.data_ptr(), v.data_ptr(), 0, n, L.X_UPDATE
    for i in range(world):
        e.weight[i] = 1.0 / world
    e.opt = L.make_optim(OptimConfig(), 1, 1e-3)
    c = emu.comm()
    if topo == "allreduce":
        L.check(L.lib().ldnn_allreduce_fused(C.byref(c), C.byref(e), None))
    else:
        L.check(L.lib().ldnn_ring_fused(C.byref(c), C.byref(e), 1 if topo == "ring" else 2, None))
    emu.ok()
print("ok")
