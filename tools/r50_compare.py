"""ResNet-50 @224 batch 32: NativeEngine (bf16 tcgen05 kernels) vs torch bf16 autocast + channels_last (cuDNN)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from ldnn_b200.engine import OptimConfig
from ldnn_b200.models import SpecNet, resnet50_spec
from ldnn_b200.models.native import NativeEngine
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
x = torch.randn(B, 3, 224, 224, device="cuda"); y = torch.randint(0, 1000, (B,), device="cuda")
def timeit(fn, n=20, w=5):
    for _ in range(w): fn()
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize(); return a.elapsed_time(b) / n
m = SpecNet(resnet50_spec()).cuda().to(memory_format=torch.channels_last)
opt = torch.optim.Adam(m.parameters(), lr=1e-3, fused=True)
xc = x.contiguous(memory_format=torch.channels_last)
def torch_step():
    opt.zero_grad(set_to_none=True)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        loss = F.cross_entropy(m(xc).float(), y)
    loss.backward(); opt.step()
t_torch = timeit(torch_step)
print(f"torch bf16 autocast channels_last: {t_torch:.2f} ms/step  {B / t_torch * 1e3:.0f} samples/s")
del m, opt
eng = NativeEngine(resnet50_spec(), "cuda:0", OptimConfig(), batch_size=B)
t_nat = timeit(lambda: eng.train_step(x, y))
print(f"native engine:                     {t_nat:.2f} ms/step  {B / t_nat * 1e3:.0f} samples/s   kernels/step {eng.launches_per_step(B)}")
