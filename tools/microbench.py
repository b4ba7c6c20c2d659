"""Steady-state per-kernel timings (CUDA events, warm L2) for the repo model's shapes at batch 64."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from ldnn_b200.ops import native_ops as K
from gpu_util import make_conv_case

def timeit(fn, iters=50):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3

def graph_timeit(fn, iters=50):
    """Same, but replayed from a CUDA graph of `iters` launches (no CPU launch overhead)."""
    for _ in range(3): fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters): fn()
    g.replay(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); g.replay(); b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3

which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which in ("all", "bn"):
    print("== BN (us, graph-replayed)")
    for M, C in [(65536, 64), (16384, 128), (4096, 256), (1024, 512), (256, 1024)]:
        dev = "cuda"
        y = torch.randn(M, C, device=dev).bfloat16(); dout = torch.randn(M, C, device=dev).bfloat16(); out = torch.empty_like(y)
        gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
        rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev); nbt = torch.zeros(1, dtype=torch.int64, device=dev)
        chunks = max(1, min(M // 128, max(1, 296 // max(1, C // 64))))
        partial = torch.zeros(chunks * 2 * C, device=dev); ticket = torch.zeros(2, dtype=torch.int32, device=dev)
        sm, si, sc, sh = (torch.zeros(C, device=dev) for _ in range(4)); dg, db = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
        dyr = torch.empty_like(y)
        t1 = graph_timeit(lambda: K.bn_fwd_stats(y, M, C, chunks, partial, ticket, gamma, beta, rm, rv, nbt, sm, si, sc, sh))
        t2 = graph_timeit(lambda: K.bn_fwd_apply(y, sc, sh, None, out, M, C, True))
        sums = torch.zeros(2 * C, device=dev)
        t2 = graph_timeit(lambda: K.bn_fwd_apply_sums(y, sums, gamma, beta, rm, rv, nbt, sm, si, None, out, M, C, True, True))
        t3 = graph_timeit(lambda: K.bn_bwd_reduce_sums(dout, out, y, sm, si, sums, M, C, chunks, True))
        t4 = graph_timeit(lambda: K.bn_bwd_apply_sums(dout, out, y, sm, si, gamma, sums, dyr, None, dg, db, M, C, True, False))
        print(f"M={M:6d} C={C:5d} chunks={chunks:4d} stats {t1:6.1f} apply {t2:6.1f} bwd_reduce {t3:6.1f} bwd_apply {t4:6.1f}")
if which in ("all", "conv"):
    print("== conv (us, graph-replayed): fprop dgrad wgrad   [TFLOP/s of fprop]")
    cases = [(64, 32, 32, 64, 64, 1, 1, 0), (64, 32, 32, 64, 128, 3, 2, 1), (64, 16, 16, 128, 128, 3, 1, 1), (64, 32, 32, 64, 128, 1, 2, 0),
             (64, 16, 16, 128, 256, 3, 2, 1), (64, 8, 8, 256, 256, 3, 1, 1), (64, 8, 8, 256, 512, 3, 2, 1), (64, 4, 4, 512, 512, 3, 1, 1),
             (64, 4, 4, 512, 1024, 3, 2, 1), (64, 2, 2, 1024, 1024, 3, 1, 1)]
    for case in cases:
        N, H, W, Cin, Cout, R, s, p = case
        x, w, dy, P, Q = make_conv_case(*case)
        y = torch.empty(N, P, Q, Cout, device="cuda", dtype=torch.bfloat16)
        dx = torch.zeros(N, H, W, Cin, device="cuda", dtype=torch.bfloat16)
        dw = torch.empty(Cout, R * R * Cin, device="cuda", dtype=torch.float32)
        pf = K.ConvPlan(K.FPROP, N, H, W, Cin, Cout, R, s, p, x, w.view(Cout, -1), y)
        pd = K.ConvPlan(K.DGRAD, N, H, W, Cin, Cout, R, s, p, dx, w.view(Cout, -1), dy, addend=dx, accumulate=True)
        pw = K.ConvPlan(K.WGRAD, N, H, W, Cin, Cout, R, s, p, x, None, dy, dw=dw, atomic_out=True)
        tf, td, tw = graph_timeit(pf.run), graph_timeit(pd.run), graph_timeit(pw.run)
        fl = 2.0 * N * P * Q * Cout * R * R * Cin
        print(f"{str(case):40s} {tf:6.1f} {td:6.1f} {tw:6.1f}   [{fl / tf / 1e6:6.1f}]  launches f/d/w {pf.num_launches}/{pd.num_launches}/{pw.num_launches}")
if which in ("all", "misc"):
    n = 44_600_000 // 4 * 4
    p_, g_, m_, v_ = (torch.zeros(n, device="cuda") for _ in range(4)); sh_ = torch.zeros(n, device="cuda", dtype=torch.bfloat16)
    from ldnn_b200.engine import OptimConfig
    from ldnn_b200.ops import _lib as L
    o = L.make_optim(OptimConfig(), 1, 1e-3)
    t = timeit(lambda: K.optim_step(p_, g_, m_, v_, sh_, n, o), 20)
    print(f"adam over {n} elems: {t:.1f} us  ({(7 * 4 + 2) * n / t / 1e6:.2f} TB/s)")
