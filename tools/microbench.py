"""Steady-state per-kernel timings for the repo model's shapes at batch 64, both precisions, with the clock record and
the roofline fraction in the same output (python tools/microbench.py > profiles/microbench_kernels_r2.txt).

Timing: CUDA events around a CUDA graph of 50 back-to-back launches (no CPU launch cost); the operands of one launch
(<= 40 MB) fit the 126 MB L2, so these are WARM-L2 numbers — an upper bound on what a layer can reach inside the step,
where its inputs were just written by the previous kernel.  Roofline of a conv = the slower of FLOPs at the measured
tensor rate (bf16: cuBLAS sustained from MEASURED_PEAKS.json; tf32: half of it, the hardware ratio) and compulsory
bytes (x + w + y) at the measured HBM copy rate."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from ldnn_b200.ops import native_ops as K
from ldnn_b200.utils.timing import ClockSampler
from gpu_util import make_conv_case

try:
    PEAKS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
except OSError:
    PEAKS = {"hbm_gbs": 6554.6, "bf16_tflops_sustained": 1363.5}
HBM, TF = PEAKS["hbm_gbs"] * 1e9, {"bf16": PEAKS["bf16_tflops_sustained"] * 1e12, "tf32": PEAKS["bf16_tflops_sustained"] * 0.5e12}
DT = {"bf16": torch.bfloat16, "tf32": torch.float32}
CASES = [(64, 32, 32, 64, 64, 1, 1, 0), (64, 32, 32, 64, 128, 3, 2, 1), (64, 16, 16, 128, 128, 3, 1, 1),
         (64, 32, 32, 64, 128, 1, 2, 0), (64, 16, 16, 128, 256, 3, 2, 1), (64, 8, 8, 256, 256, 3, 1, 1),
         (64, 8, 8, 256, 512, 3, 2, 1), (64, 4, 4, 512, 512, 3, 1, 1), (64, 4, 4, 512, 1024, 3, 2, 1),
         (64, 2, 2, 1024, 1024, 3, 1, 1)]


def graph_timeit(fn, iters=50):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters):
            fn()
    g.replay(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); g.replay(); b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3


def tuned(mode, case, dcode, mk):
    bn, sp, var = K.tuned_conv_config(mode, case, None, enabled=False, dtype=dcode)    # shipped tunings, no timing runs
    return mk(block_n=bn, splits=sp, **K.variant_kwargs(var))


def convs(dt):
    dtype, dcode, esz = DT[dt], (1 if dt == "tf32" else 0), (4 if dt == "tf32" else 2)
    print(f"== conv {dt} (us, graph-replayed, shipped tunings): fprop dgrad wgrad | roofline us | fractions")
    tot, roof_tot = 0.0, 0.0
    for case in CASES:
        N, H, W, Cin, Cout, R, s, p = case
        x, w, dy, P, Q = make_conv_case(*case, dtype=dtype)
        y = torch.empty(N, P, Q, Cout, device="cuda", dtype=dtype)
        dx = torch.zeros(N, H, W, Cin, device="cuda", dtype=dtype)
        dw = torch.zeros(Cout, R * R * Cin, device="cuda", dtype=torch.float32)
        pf = tuned(K.FPROP, case, dcode, lambda **kw: K.ConvPlan(K.FPROP, *case, x, w.view(Cout, -1), y, **kw))
        pd = tuned(K.DGRAD, case, dcode, lambda **kw: K.ConvPlan(K.DGRAD, *case, dx, w.view(Cout, -1), dy, addend=dx,
                                                                   accumulate=True, **kw))
        pw = tuned(K.WGRAD, case, dcode, lambda **kw: K.ConvPlan(K.WGRAD, *case, x, None, dy, dw=dw, atomic_out=True, **kw))
        ts = [graph_timeit(pf.run), graph_timeit(pd.run), graph_timeit(pw.run)]
        fl = 2.0 * N * P * Q * Cout * R * R * Cin
        byts = esz * (N * H * W * Cin + Cout * Cin * R * R + N * P * Q * Cout)
        roof = max(fl / TF[dt], byts / HBM) * 1e6
        tot += sum(ts); roof_tot += 3 * roof
        print(f"{str(case):36s} {ts[0]:6.1f} {ts[1]:6.1f} {ts[2]:6.1f} | {roof:5.1f} | "
              + " ".join(f"{roof / t:4.2f}" for t in ts) + f"   grids {pf.grid} {pd.grid} {pw.grid}")
    print(f"sum {tot:.0f} us vs roofline {roof_tot:.0f} us -> {roof_tot / tot:.2f}")


def bn(dt):
    dtype, esz = DT[dt], (4 if dt == "tf32" else 2)
    print(f"== BN {dt} (us, graph-replayed): fwd apply(+stats from sums) | bwd reduce | bwd apply | HBM roofline us of each")
    for M, C in [(65536, 64), (16384, 128), (4096, 256), (1024, 512), (256, 1024)]:
        dev = "cuda"
        y = torch.randn(M, C, device=dev).to(dtype); dout = torch.randn(M, C, device=dev).to(dtype); out = torch.empty_like(y)
        gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
        rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev); nbt = torch.zeros(1, dtype=torch.int64, device=dev)
        chunks = max(1, min(M // 128, max(1, 296 // max(1, C // 64))))
        sm, si = torch.zeros(C, device=dev), torch.ones(C, device=dev)
        dg, db = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
        dyr, sums = torch.empty_like(y), torch.zeros(2 * C, device=dev)
        t2 = graph_timeit(lambda: K.bn_fwd_apply_sums(y, sums, gamma, beta, rm, rv, nbt, sm, si, None, out, M, C, True, True))
        t3 = graph_timeit(lambda: K.bn_bwd_reduce_sums(dout, out, y, sm, si, sums, M, C, chunks, True))
        t4 = graph_timeit(lambda: K.bn_bwd_apply_sums(dout, out, y, sm, si, gamma, sums, dyr, None, dg, db, M, C, True, False))
        r2, r3, r4 = (2 * M * C * esz / HBM * 1e6), (3 * M * C * esz / HBM * 1e6), (4 * M * C * esz / HBM * 1e6)
        print(f"M={M:6d} C={C:5d}  {t2:6.1f} | {t3:6.1f} | {t4:6.1f} |  {r2:4.1f} {r3:4.1f} {r4:4.1f}")


def misc():
    n = 44_600_000 // 4 * 4
    p_, g_, m_, v_ = (torch.zeros(n, device="cuda") for _ in range(4))
    sh_ = torch.zeros(n, device="cuda", dtype=torch.bfloat16)
    from ldnn_b200.engine import OptimConfig
    from ldnn_b200.ops import _lib as L
    o = L.make_optim(OptimConfig(), 1, 1e-3)
    for name, sh, byts in (("tf32 engine (no shadow)", None, 7 * 4), ("bf16 engine (+bf16 shadow)", sh_, 7 * 4 + 2)):
        t = graph_timeit(lambda: K.optim_step(p_, g_, m_, v_, sh, n, o), 10)
        print(f"flat Adam, {name}: {n} elems in {t:.1f} us = {byts * n / t / 1e6:.2f} TB/s = "
              f"{byts * n / t * 1e6 / HBM:.2f} of the HBM copy rate")


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    with ClockSampler(0, period_s=0.1) as clk:
        for dt in ("tf32", "bf16"):
            if which in ("all", "conv"):
                convs(dt)
            if which in ("all", "bn"):
                bn(dt)
        if which in ("all", "misc"):
            misc()
    print(f"# denominators: HBM copy {HBM / 1e9:.0f} GB/s, tensor bf16 {TF['bf16'] / 1e12:.0f} / tf32 {TF['tf32'] / 1e12:.0f} TFLOP/s (of measured)")
    print(f"# clocks during these measurements: {json.dumps(clk.summary())}")
