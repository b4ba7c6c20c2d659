"""Accuracy of the kind::tf32 conv kernels against an fp64 oracle, for both TMA element types
(LDNN_TF32_TMA=float32: raw fp32 bits, the MMA truncates; tfloat32: the TMA unit converts), next to cuDNN's own
TF32 convolution on the same tensors.  Prints one line per (shape, mode)."""
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from ldnn_b200.ops import native_ops as K  # noqa: E402

CASES = [(64, 16, 16, 128, 128, 3, 1, 1), (64, 8, 8, 256, 512, 3, 2, 1), (64, 2, 2, 1024, 1024, 3, 1, 1)]


def rel(a, b):
    return float((a.double() - b).abs().max() / b.abs().max())


def rms(a, b):
    return float(((a.double() - b) ** 2).mean().sqrt() / (b ** 2).mean().sqrt())


for case in CASES:
    N, H, W, Cin, Cout, R, s, p = case
    g = torch.Generator().manual_seed(0)
    P, Q = (H + 2 * p - R) // s + 1, (W + 2 * p - R) // s + 1
    x = (torch.randn(N, H, W, Cin, generator=g) * 0.5).cuda()
    w = (torch.randn(Cout, R, R, Cin, generator=g) / (R * R * Cin) ** 0.5).cuda()
    dy = (torch.randn(N, P, Q, Cout, generator=g) * 0.5).cuda()
    xd = x.double().permute(0, 3, 1, 2).requires_grad_(True)
    wd = w.double().permute(0, 3, 1, 2).requires_grad_(True)
    yd = F.conv2d(xd, wd, None, s, p)
    yd.backward(dy.double().permute(0, 3, 1, 2))
    y64, dx64, dw64 = yd.detach().permute(0, 2, 3, 1), xd.grad.permute(0, 2, 3, 1), wd.grad.permute(0, 2, 3, 1)
    torch.backends.cudnn.allow_tf32 = True
    xc = x.permute(0, 3, 1, 2).clone().requires_grad_(True)
    wc = w.permute(0, 3, 1, 2).clone().requires_grad_(True)
    yc = F.conv2d(xc, wc, None, s, p)
    yc.backward(dy.permute(0, 3, 1, 2))
    print(f"{case} cudnn-tf32  fprop max {rel(yc.detach().permute(0,2,3,1), y64):.2e} rms {rms(yc.detach().permute(0,2,3,1), y64):.2e} | "
          f"dgrad max {rel(xc.grad.permute(0,2,3,1), dx64):.2e} | wgrad max {rel(wc.grad.permute(0,2,3,1), dw64):.2e}")
    y = torch.empty(N, P, Q, Cout, device="cuda")
    K.ConvPlan(K.FPROP, *case, x, w.view(Cout, -1), y).run()
    dx = torch.zeros(N, H, W, Cin, device="cuda")
    K.ConvPlan(K.DGRAD, *case, dx, w.view(Cout, -1), dy, addend=dx, accumulate=True).run()
    dw = torch.zeros(Cout, R * R * Cin, device="cuda")
    K.ConvPlan(K.WGRAD, *case, x, None, dy, dw=dw).run()
    torch.cuda.synchronize()
    print(f"{case} ours[{os.environ.get('LDNN_TF32_TMA', 'tfloat32')}] fprop max {rel(y, y64):.2e} rms {rms(y, y64):.2e} | "
          f"dgrad max {rel(dx, dx64):.2e} | wgrad max {rel(dw.view(Cout, R, R, Cin), dw64):.2e}")
