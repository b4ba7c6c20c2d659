"""tcgen05 implicit-GEMM conv (fprop / dgrad / wgrad) vs fp32 torch conv, at every shape of the
repo model plus ResNet-50-style shapes."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

# (N, H, W, Cin, Cout, R, stride, pad)
ENHANCED = [
    (64, 32, 32, 64, 128, 3, 2, 1), (64, 16, 16, 128, 128, 3, 1, 1), (64, 32, 32, 64, 128, 1, 2, 0),
    (64, 16, 16, 128, 256, 3, 2, 1), (64, 8, 8, 256, 256, 3, 1, 1), (64, 16, 16, 128, 256, 1, 2, 0),
    (64, 8, 8, 256, 512, 3, 2, 1), (64, 4, 4, 512, 512, 3, 1, 1), (64, 8, 8, 256, 512, 1, 2, 0),
    (64, 4, 4, 512, 1024, 3, 2, 1), (64, 2, 2, 1024, 1024, 3, 1, 1), (64, 4, 4, 512, 1024, 1, 2, 0),
    (64, 32, 32, 64, 64, 1, 1, 0),      # the stem after im2col (K padded to 64)
]
ODD = [
    (6, 16, 16, 128, 128, 3, 1, 1),     # partial batch (upstream keeps the last short batch)
    (5, 2, 2, 1024, 1024, 3, 1, 1),
    (3, 14, 14, 256, 256, 3, 1, 1),     # non power-of-two image (ResNet-50 stage 4)
    (2, 28, 28, 128, 128, 3, 2, 1),
    (2, 56, 56, 64, 256, 1, 1, 0),
    (2, 7, 7, 512, 2048, 1, 1, 0),
]


DT = ["bf16", "tf32"]


def _run(case, mode, dt="bf16", **kw):
    """max-rel error of one native conv launch vs the fp32 cuDNN reference (TF32 off).  dt = "tf32": fp32
    tensors in, tcgen05 kind::tf32 (the reference's precision); "bf16": bf16 tensors, kind::f16."""
    from gpu_util import DTYPES, fp32_mode, make_conv_case, ref_conv, ref_conv_grads, rel_err
    from ldnn_b200.ops import native_ops as K
    fp32_mode()
    N, H, W, Cin, Cout, R, s, p = case
    dtype = DTYPES[dt]
    x, w, dy, P, Q = make_conv_case(*case, dtype=dtype)
    if mode == K.FPROP:
        y = torch.full((N, P, Q, Cout), float("nan"), device="cuda", dtype=dtype)
        K.ConvPlan(K.FPROP, N, H, W, Cin, Cout, R, s, p, x, w.view(Cout, -1), y, **kw).run()
        torch.cuda.synchronize()
        return rel_err(y, ref_conv(x, w, s, p))
    dx_ref, dw_ref = ref_conv_grads(x, w, dy, s, p)
    if mode == K.DGRAD:
        dx = torch.zeros(N, H, W, Cin, device="cuda", dtype=dtype)
        K.ConvPlan(K.DGRAD, N, H, W, Cin, Cout, R, s, p, dx, w.view(Cout, -1), dy, addend=dx, accumulate=True, **kw).run()
        torch.cuda.synchronize()
        return rel_err(dx, dx_ref)
    fill = 0.0 if kw.get("atomic_out") else float("nan")
    dw = torch.full((Cout, R * R * Cin), fill, device="cuda", dtype=torch.float32)
    K.ConvPlan(K.WGRAD, N, H, W, Cin, Cout, R, s, p, x, None, dy, dw=dw, **kw).run()
    torch.cuda.synchronize()
    return rel_err(dw.view(Cout, R, R, Cin), dw_ref)


def _tol(dt):
    from gpu_util import TOL
    return TOL[dt]


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("case", ENHANCED + ODD)
def test_fprop(case, dt):
    from ldnn_b200.ops import native_ops as K
    assert _run(case, K.FPROP, dt) < _tol(dt)[0]


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("case", [c for c in ENHANCED + ODD])
def test_dgrad(case, dt):
    from ldnn_b200.ops import native_ops as K
    assert _run(case, K.DGRAD, dt) < _tol(dt)[0]


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("case", ENHANCED + ODD)
def test_wgrad(case, dt):
    from ldnn_b200.ops import native_ops as K
    assert _run(case, K.WGRAD, dt) < _tol(dt)[1]


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("case", ENHANCED[:6] + ODD[:3])
def test_wgrad_atomic_split_k(case, dt):
    """split-K partial tiles red.add'ed into a pre-zeroed dw (the engine's fast path)."""
    from ldnn_b200.ops import native_ops as K
    assert _run(case, K.WGRAD, dt, atomic_out=True) < _tol(dt)[1]


@pytest.mark.parametrize("dt", DT)
@pytest.mark.parametrize("splits", [1, 2, 5])
@pytest.mark.parametrize("block_n", [64, 128, 256])
def test_tile_and_split_variants(splits, block_n, dt):
    from ldnn_b200.ops import native_ops as K
    case = (8, 8, 8, 256, 256, 3, 1, 1)
    assert _run(case, K.FPROP, dt, splits=splits, block_n=block_n) < _tol(dt)[0]
    assert _run(case, K.DGRAD, dt, splits=splits, block_n=block_n) < _tol(dt)[0]
    assert _run(case, K.WGRAD, dt, splits=splits, block_n=block_n) < _tol(dt)[1]


@pytest.mark.parametrize("dt", DT)
def test_split_k_is_deterministic_and_reusable(dt):
    from gpu_util import DTYPES, make_conv_case
    from ldnn_b200.ops import native_ops as K
    N, H, W, Cin, Cout, R, s, p = case = (64, 2, 2, 1024, 1024, 3, 1, 1)
    x, w, dy, P, Q = make_conv_case(*case, dtype=DTYPES[dt])
    y = torch.empty(N, P, Q, Cout, device="cuda", dtype=DTYPES[dt])
    plan = K.ConvPlan(K.FPROP, N, H, W, Cin, Cout, R, s, p, x, w.view(Cout, -1), y)
    assert plan.scratch_bytes > 0
    plan.run(); torch.cuda.synchronize(); y0 = y.clone()
    for _ in range(3):
        y.zero_(); plan.run(); torch.cuda.synchronize()
        assert torch.equal(y, y0)


@pytest.mark.parametrize("case", [(64, 16, 16, 128, 128, 3, 1, 1), (6, 8, 8, 256, 512, 3, 2, 1), (64, 2, 2, 1024, 1024, 3, 1, 1)])
@pytest.mark.parametrize("relu", [True, False])
@pytest.mark.parametrize("dt", DT)
def test_dgrad_fused_bn_backward_reduce(case, relu, dt):
    """dgrad epilogue also accumulates (sum dy, sum dy*xhat) of the layer that owns dx."""
    from gpu_util import DTYPES, make_conv_case
    from ldnn_b200.ops import native_ops as K
    N, H, W, Cin, Cout, R, s, p = case
    dtype = DTYPES[dt]
    x, w, dy, P, Q = make_conv_case(*case, dtype=dtype)
    g = torch.Generator().manual_seed(11)
    out = torch.randn(N, H, W, Cin, generator=g).cuda().to(dtype)          # owner's post-activation output
    y = torch.randn(N, H, W, Cin, generator=g).cuda().to(dtype)            # owner's pre-BN output
    mean, invstd = torch.randn(Cin, generator=g).cuda() * 0.1, (torch.rand(Cin, generator=g) + 0.5).cuda()
    sums = torch.zeros(2 * Cin, device="cuda")
    dx = torch.zeros(N, H, W, Cin, device="cuda", dtype=dtype)
    K.ConvPlan(K.DGRAD, N, H, W, Cin, Cout, R, s, p, dx, w.view(Cout, -1), dy, addend=dx, accumulate=True,
               bn_bwd_fuse=(out, y, mean, invstd, sums, relu)).run()
    torch.cuda.synchronize()
    d = dx.float().view(-1, Cin)
    if relu:
        d = d * (out.float().view(-1, Cin) > 0)
    xh = (y.float().view(-1, Cin) - mean) * invstd
    want_a, want_b = d.sum(0), (d * xh).sum(0)
    assert torch.allclose(sums[:Cin], want_a, rtol=2e-3, atol=2e-3 * float(want_a.abs().max()))
    assert torch.allclose(sums[Cin:], want_b, rtol=2e-3, atol=2e-3 * float(want_b.abs().max()))


@pytest.mark.parametrize("case", [(64, 16, 16, 128, 128, 3, 1, 1), (64, 8, 8, 256, 256, 3, 1, 1),
                                  (64, 16, 16, 128, 256, 3, 2, 1), (8, 14, 14, 256, 256, 3, 1, 1)])
@pytest.mark.parametrize("splits", [1, 3])
@pytest.mark.parametrize("dt", DT)
def test_cluster_multicast_weights(case, splits, dt):
    """Clusters of two M tiles: each CTA loads half of the weight tile and TMA-multicasts it to both."""
    from ldnn_b200.ops import native_ops as K
    assert _run(case, K.FPROP, dt, cluster=2, block_n=128, splits=splits) < _tol(dt)[0]
    assert _run(case, K.DGRAD, dt, cluster=2, block_n=128, splits=splits) < _tol(dt)[0]


@pytest.mark.parametrize("case", [(64, 16, 16, 128, 128, 3, 1, 1), (64, 8, 8, 256, 256, 3, 1, 1),
                                  (64, 16, 16, 128, 256, 3, 2, 1), (8, 14, 14, 256, 256, 3, 1, 1)])
@pytest.mark.parametrize("block_n", [128, 256])
@pytest.mark.parametrize("splits", [1, 3])
@pytest.mark.parametrize("dt", DT)
def test_cta_pair_umma_m256(case, block_n, splits, dt):
    """cluster=-2: the two CTAs of a pair run one tcgen05.mma.cta_group::2 (UMMA_M = 256); each loads half of B."""
    from ldnn_b200.ops import native_ops as K
    N, H, W, Cin, Cout, R, s, p = case
    if Cout % block_n == 0:
        assert _run(case, K.FPROP, dt, cluster=-2, block_n=block_n, splits=splits) < _tol(dt)[0]
    if Cin % block_n == 0:
        assert _run(case, K.DGRAD, dt, cluster=-2, block_n=block_n, splits=splits) < _tol(dt)[0]


@pytest.mark.parametrize("case", [(64, 16, 16, 128, 128, 3, 1, 1), (8, 56, 56, 64, 256, 1, 1, 0), (6, 8, 8, 256, 512, 3, 2, 1)])
@pytest.mark.parametrize("block_n", [64, 128, 256])
@pytest.mark.parametrize("dt", DT)
def test_lite_launch_two_ctas_per_sm(case, block_n, dt):
    """Short-pipeline launch (fewer smem stages, 4 epilogue warps) used for large grids."""
    from ldnn_b200.ops import native_ops as K
    N, H, W, Cin, Cout, R, s, p = case
    if Cout % block_n == 0:
        assert _run(case, K.FPROP, dt, lite=True, block_n=block_n) < _tol(dt)[0]
    if Cin % block_n == 0:
        assert _run(case, K.DGRAD, dt, lite=True, block_n=block_n) < _tol(dt)[0]
        assert _run(case, K.WGRAD, dt, lite=True, block_n=block_n, atomic_out=True) < _tol(dt)[1]
