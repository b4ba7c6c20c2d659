"""Fused BN(+residual,+ReLU) fwd/bwd, the pool+fc+CE head and the flat optimizer vs fp32 torch."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

# activation storage types of the native kernels and how far their outputs may sit from the fp32 torch oracle:
# bf16 rounds every stored activation (2^-9 relative), fp32 storage (the TF32 path) only differs by summation order
DT = {"bf16": torch.bfloat16, "tf32": torch.float32}
ACT_TOL = {"bf16": dict(atol=3e-2, rtol=2e-2), "tf32": dict(atol=2e-4, rtol=2e-4)}
GRAD_TOL = {"bf16": 3e-2, "tf32": 5e-4}


@pytest.mark.parametrize("M,C", [(64 * 32 * 32, 64), (64 * 16 * 16, 128), (64 * 2 * 2, 1024), (6 * 8 * 8, 256),
                                 (37, 2048), (100, 192), (64 * 4 * 4, 512)])
@pytest.mark.parametrize("relu,use_res", [(True, True), (True, False), (False, False)])
@pytest.mark.parametrize("dt", list(DT))
def test_bn_forward_backward(M, C, relu, use_res, dt):
    from ldnn_b200.ops import native_ops as K
    dev = "cuda"
    adt = DT[dt]
    g = torch.Generator().manual_seed(M + C)
    y = (torch.randn(M, C, generator=g) * 2 + 0.5).to(dev).to(adt)
    res = torch.randn(M, C, generator=g).to(dev).to(adt) if use_res else None
    dout = torch.randn(M, C, generator=g).to(dev).to(adt)
    gamma = (torch.rand(C, generator=g) + 0.5).to(dev)
    beta = torch.randn(C, generator=g).to(dev)
    rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    nbt = torch.zeros(1, dtype=torch.int64, device=dev)
    chunks = max(1, min(296, M // 32))
    partial = torch.zeros(chunks * 2 * C, device=dev)
    ticket = torch.zeros(2, dtype=torch.int32, device=dev)
    sm, si, sc, sh = (torch.zeros(C, device=dev) for _ in range(4))
    out = torch.empty(M, C, device=dev, dtype=adt)
    for _ in range(2):        # twice: tickets must self-reset
        K.bn_fwd_stats(y, M, C, chunks, partial, ticket, gamma, beta, rm, rv, nbt, sm, si, sc, sh)
        K.bn_fwd_apply(y, sc, sh, res, out, M, C, relu)
    # reference (fp32 on the same bf16 inputs)
    yr = y.float().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    rrm, rrv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    for _ in range(2):
        o = F.batch_norm(yr, rrm, rrv, gr, br, True, 0.1, 1e-5)
    rr = res.float().requires_grad_(True) if use_res else None
    if use_res:
        o = o + rr
    if relu:
        o = F.relu(o)
    assert torch.allclose(out.float(), o, **ACT_TOL[dt])
    assert torch.allclose(rm, rrm, atol=1e-4) and torch.allclose(rv, rrv, atol=1e-3, rtol=1e-3)
    assert int(nbt) == 2
    # backward; the kernel masks with its own bf16 output, mirror that in the reference
    dgamma, dbeta = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
    dyr = torch.empty(M, C, device=dev, dtype=adt)
    dres = torch.empty(M, C, device=dev, dtype=adt) if use_res else None
    K.bn_bwd_reduce(dout, out, y, sm, si, partial, ticket[1:], M, C, chunks, relu, dgamma, dbeta)
    K.bn_bwd_apply(dout, out, y, sm, si, gamma, dgamma, dbeta, dyr, dres, M, C, relu, False)
    o.backward(dout.float())
    scale = float(yr.grad.abs().max()) + 1e-6
    gt_ = GRAD_TOL[dt]
    bad = ((dyr.float() - yr.grad).abs() / scale > gt_).float().mean()      # ReLU-mask flips at ~0, see below
    assert float(bad) < 2e-3, float(bad)
    assert torch.allclose(dgamma, gr.grad, rtol=gt_, atol=gt_ * float(gr.grad.abs().max()))
    assert torch.allclose(dbeta, br.grad, rtol=gt_, atol=gt_ * float(br.grad.abs().max()))
    if use_res:
        bad = ((dres.float() - rr.grad).abs() > 2e-2 + 2e-2 * rr.grad.abs()).float().mean()   # mask flips at ~0
        assert float(bad) < 2e-3, float(bad)


def test_bn_eval_coeffs():
    from ldnn_b200.ops import native_ops as K
    C = 256
    g, b, rm = torch.rand(C, device="cuda") + 0.5, torch.randn(C, device="cuda"), torch.randn(C, device="cuda")
    rv = torch.rand(C, device="cuda") + 0.1
    sc, sh = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")
    K.bn_eval_coeffs(g, b, rm, rv, sc, sh, C)
    y = torch.randn(50, C, device="cuda").bfloat16()
    out = torch.empty_like(y)
    K.bn_fwd_apply(y, sc, sh, None, out, 50, C, False)
    ref = F.batch_norm(y.float(), rm, rv, g, b, False, 0.1, 1e-5)
    assert torch.allclose(out.float(), ref, atol=3e-2, rtol=2e-2)


@pytest.mark.parametrize("N,HW,C,classes", [(64, 4, 1024, 10), (7, 4, 1024, 10), (16, 49, 2048, 1000)])
@pytest.mark.parametrize("dt", list(DT))
def test_head(N, HW, C, classes, dt):
    from ldnn_b200.ops import native_ops as K
    dev = "cuda"
    g = torch.Generator().manual_seed(N)
    act = torch.randn(N, HW, C, generator=g).to(dev).to(DT[dt])
    fw = (torch.randn(classes, C, generator=g) / math.sqrt(C)).to(dev)
    fb = torch.randn(classes, generator=g).to(dev) * 0.1
    labels = torch.randint(0, classes, (N,), generator=g).to(dev)
    f = lambda *s: torch.zeros(*s, device=dev)
    pooled, logits, dlogits, loss_n, correct_n = f(N, C), f(N, classes), f(N, classes), f(N), f(N)
    pred = torch.zeros(N, dtype=torch.int64, device=dev)
    dact = torch.zeros(N, HW, C, device=dev, dtype=DT[dt])
    dfw, dfb, loss, correct = f(classes, C), f(classes), f(1), f(1)
    K.head_fwd_bwd(act, N, HW, C, classes, fw, fb, labels, pooled, logits, dlogits, loss_n, correct_n, pred, dact,
                   dfw, dfb, loss, correct, True)
    a = act.float().requires_grad_(True)
    w, b = fw.clone().requires_grad_(True), fb.clone().requires_grad_(True)
    lg = F.linear(a.mean(1), w, b)
    l = F.cross_entropy(lg, labels)
    l.backward()
    assert torch.allclose(logits, lg, atol=1e-3, rtol=1e-3)
    assert abs(float(loss) - float(l)) < 1e-3
    assert int(correct) == int((lg.argmax(1) == labels).sum())
    assert torch.equal(pred, lg.argmax(1))
    assert torch.allclose(dfw, w.grad, atol=1e-4, rtol=1e-3) and torch.allclose(dfb, b.grad, atol=1e-5, rtol=1e-3)
    assert float((dact.float() - a.grad).abs().max()) <= (1e-2 if dt == "bf16" else 1e-5) * float(a.grad.abs().max()) + 1e-8


@pytest.mark.parametrize("kind", ["adam", "sgd"])
def test_flat_optimizer_kernel(kind):
    from ldnn_b200.engine import FlatState, OptimConfig, Slot
    from ldnn_b200.ops import _lib as L, native_ops as K
    n = 4096 * 33
    opt = OptimConfig(kind=kind, lr=3e-3, weight_decay=1e-2)
    ref = FlatState([Slot("w", (n,), 0, n)], n, "cuda", opt)
    ref.params.normal_()
    p, m, v = ref.params.clone(), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    shadow = torch.zeros(n, device="cuda", dtype=torch.bfloat16)
    for step in range(1, 6):
        g = torch.randn(n, device="cuda")
        ref.optimizer_step(g)
        K.optim_step(p, g, m, v if kind == "adam" else None, shadow, n, L.make_optim(opt, step, ref.lr))
    assert torch.allclose(p, ref.params, atol=1e-6, rtol=1e-5)
    assert torch.equal(shadow, p.bfloat16())


@pytest.mark.parametrize("dt", list(DT))
def test_im2col_stem_matches_unfold(dt):
    from ldnn_b200.ops import native_ops as K
    N, C, H, W, R = 5, 3, 32, 32, 3
    x = torch.randn(N, C, H, W, device="cuda")
    col = torch.empty(N * H * W, 64, device="cuda", dtype=DT[dt])
    K.im2col_stem(x, col, N, C, H, W, R, 1, 1, 64)
    ref = F.unfold(x, R, padding=1).view(N, C, R * R, H * W).permute(0, 3, 2, 1).reshape(N * H * W, R * R * C)
    assert torch.equal(col[:, :27], ref.to(DT[dt])) and float(col[:, 27:].abs().max()) == 0.0


@pytest.mark.parametrize("case", [(64, 16, 16, 128, 128, 3, 1, 1), (64, 2, 2, 1024, 1024, 3, 1, 1), (6, 8, 8, 256, 512, 3, 2, 1),
                                  (3, 14, 14, 256, 256, 3, 1, 1)])
@pytest.mark.parametrize("use_res", [False, True])
@pytest.mark.parametrize("dt", list(DT))
def test_conv_fused_stats_and_sums_bn(case, use_res, dt):
    """conv epilogue accumulates sum / sum-of-squares; BN fwd/bwd 'sums' kernels vs fp32 torch."""
    from gpu_util import fp32_mode, make_conv_case, ref_conv
    from ldnn_b200.ops import native_ops as K
    fp32_mode()
    N, H, W, Cin, Cout, R, s, p = case
    adt = DT[dt]
    x, w, dy, P, Q = make_conv_case(*case, dtype=adt)
    M, C = N * P * Q, Cout
    y = torch.empty(N, P, Q, Cout, device="cuda", dtype=adt)
    sums = torch.zeros(2 * C, device="cuda")
    plan = K.ConvPlan(K.FPROP, N, H, W, Cin, Cout, R, s, p, x, w.view(Cout, -1), y, bn_sums=sums)
    plan.run()
    torch.cuda.synchronize()
    yr = ref_conv(x, w, s, p).view(M, C)
    assert torch.allclose(sums[:C], yr.sum(0), rtol=2e-3, atol=2e-3 * float(yr.abs().sum(0).max()))
    assert torch.allclose(sums[C:], (yr * yr).sum(0), rtol=5e-3)
    g = torch.Generator().manual_seed(5)
    gamma, beta = (torch.rand(C, generator=g) + 0.5).cuda(), torch.randn(C, generator=g).cuda()
    res = torch.randn(M, C, generator=g).cuda().to(adt) if use_res else None
    dout = torch.randn(M, C, generator=g).cuda().to(adt)
    rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    nbt = torch.zeros(1, dtype=torch.int64, device="cuda")
    sm, si = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    out = torch.empty(M, C, device="cuda", dtype=adt)
    K.bn_fwd_apply_sums(y, sums, gamma, beta, rm, rv, nbt, sm, si, res, out, M, C, True, True)
    yy = y.view(M, C).float().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    rrm, rrv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
    o = F.batch_norm(yy, rrm, rrv, gr, br, True, 0.1, 1e-5)
    rr = res.float().requires_grad_(True) if use_res else None
    o = F.relu(o + rr) if use_res else F.relu(o)
    # statistics come from the fp32 accumulators (before the store rounding), the oracle's from the stored y
    assert torch.allclose(out.float(), o, atol=4e-2 if dt == "bf16" else 2e-3, rtol=2e-2 if dt == "bf16" else 1e-3)
    assert torch.allclose(rm, rrm, atol=2e-3) and torch.allclose(rv, rrv, atol=2e-3, rtol=2e-3) and int(nbt) == 1
    bsum = torch.zeros(2 * C, device="cuda")
    dgamma, dbeta = torch.zeros(C, device="cuda"), torch.zeros(C, device="cuda")
    dyr = torch.empty(M, C, device="cuda", dtype=adt)
    dres = torch.empty(M, C, device="cuda", dtype=adt) if use_res else None
    chunks = max(1, min(M // 128, max(1, 296 // max(1, C // 64))))
    K.bn_bwd_reduce_sums(dout, out, y, sm, si, bsum, M, C, chunks, True)
    K.bn_bwd_apply_sums(dout, out, y, sm, si, gamma, bsum, dyr, dres, dgamma, dbeta, M, C, True, False)
    o.backward(dout.float())
    scale = float(yy.grad.abs().max()) + 1e-6
    # pre-activations within rounding distance of 0 may flip the ReLU mask between the bf16 kernel and
    # the fp32 oracle; allow a vanishing fraction of such elements
    gt_ = 4e-2 if dt == "bf16" else 2e-3
    bad = ((dyr.float() - yy.grad).abs() / scale > gt_).float().mean()
    assert float(bad) < 2e-3, float(bad)
    for got, want in ((dgamma, gr.grad), (dbeta, br.grad)):       # a flipped ReLU mask moves one channel's sum
        off = (got - want).abs() > gt_ * want.abs() + gt_ * float(want.abs().max())
        assert float(off.float().mean()) < 0.02, float(off.float().mean())
    if use_res:
        bad = ((dres.float() - rr.grad).abs() > 2e-2 + 2e-2 * rr.grad.abs()).float().mean()   # mask flips at ~0
        assert float(bad) < 2e-3, float(bad)
    # eval mode ignores the sums
    K.bn_fwd_apply_sums(y, None, gamma, beta, rm, rv, nbt, sm, si, None, out, M, C, False, False)
    ref = F.batch_norm(y.view(M, C).float(), rm, rv, gamma, beta, False, 0.1, 1e-5)
    assert torch.allclose(out.float(), ref, atol=4e-2, rtol=2e-2)


@pytest.mark.parametrize("N,H,W,C,k,s,p", [(4, 112, 112, 64, 3, 2, 1), (3, 9, 7, 128, 3, 2, 1), (2, 8, 8, 64, 2, 2, 0)])
@pytest.mark.parametrize("dt", list(DT))
def test_maxpool_forward_backward(N, H, W, C, k, s, p, dt):
    from ldnn_b200.ops import native_ops as K
    g = torch.Generator().manual_seed(H)
    adt = DT[dt]
    x = torch.randn(N, H, W, C, generator=g).cuda().to(adt)
    P, Q = (H + 2 * p - k) // s + 1, (W + 2 * p - k) // s + 1
    y = torch.empty(N, P, Q, C, device="cuda", dtype=adt)
    idx = torch.zeros(N * P * Q * C, device="cuda", dtype=torch.uint8)
    dy = torch.randn(N, P, Q, C, generator=g).cuda().to(adt)
    dx = torch.empty(N, H, W, C, device="cuda", dtype=adt)
    K.maxpool_fwd(x, y, idx, N, H, W, C, k, s, p)
    K.maxpool_bwd(idx, dy, dx, N, H, W, C, k, s, p)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    yr = F.max_pool2d(xr, k, s, p)
    yr.backward(dy.float().permute(0, 3, 1, 2))
    assert torch.equal(y.float(), yr.detach().permute(0, 2, 3, 1))
    assert torch.allclose(dx.float(), xr.grad.permute(0, 2, 3, 1), atol=2e-2, rtol=2e-2)
