"""Helpers shared by the GPU tests: fp32 torch references for the sm_100a kernels."""
import torch
import torch.nn.functional as F


def fp32_mode():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False


# element types of the native conv: bf16 storage (kind::f16) and fp32 storage (kind::tf32), with the max-rel
# tolerances of the output against an fp32 cuDNN reference with TF32 off: (activations, weight gradient)
DTYPES = {"bf16": torch.bfloat16, "tf32": torch.float32}
TOL = {"bf16": (1.5e-2, 5e-3), "tf32": (2e-3, 2e-3)}


def make_conv_case(N, H, W, Cin, Cout, R, stride, pad, seed=0, dev="cuda", dtype=torch.bfloat16):
    g = torch.Generator(device="cpu").manual_seed(seed)
    P, Q = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - R) // stride + 1
    x = (torch.randn(N, H, W, Cin, generator=g) * 0.5).to(dev).to(dtype)
    w = (torch.randn(Cout, R, R, Cin, generator=g) / (R * R * Cin) ** 0.5).to(dev).to(dtype)
    dy = (torch.randn(N, P, Q, Cout, generator=g) * 0.5).to(dev).to(dtype)
    return x, w, dy, P, Q


def ref_conv(x_nhwc, w_krsc, stride, pad):
    x = x_nhwc.float().permute(0, 3, 1, 2)
    w = w_krsc.float().permute(0, 3, 1, 2)
    return F.conv2d(x, w, None, stride, pad).permute(0, 2, 3, 1).contiguous()


def ref_conv_grads(x_nhwc, w_krsc, dy_nhwc, stride, pad):
    x = x_nhwc.float().permute(0, 3, 1, 2).requires_grad_(True)
    w = w_krsc.float().permute(0, 3, 1, 2).requires_grad_(True)
    y = F.conv2d(x, w, None, stride, pad)
    y.backward(dy_nhwc.float().permute(0, 3, 1, 2))
    return x.grad.permute(0, 2, 3, 1).contiguous(), w.grad.permute(0, 2, 3, 1).contiguous()


def rel_err(a, b):
    return float((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-12))
