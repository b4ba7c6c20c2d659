import os, sys
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import torch
from ldnn_b200.ops import native_ops as K
from gpu_util import make_conv_case
sys.path.insert(0, os.path.join(os.getcwd(), "tools"))
def graph_timeit(fn, iters=50):
    for _ in range(3): fn()
    torch.cuda.synchronize(); g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters): fn()
    g.replay(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); g.replay(); b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3
for case in [(64, 16, 16, 128, 128, 3, 1, 1), (64, 8, 8, 256, 256, 3, 1, 1), (64, 16, 16, 128, 256, 3, 2, 1), (64, 56, 56, 256, 256, 1, 1, 0), (64, 4, 4, 512, 512, 3, 1, 1), (64, 28, 28, 128, 128, 3, 1, 1)]:
    N, H, W, Cin, Cout, R, s, p = case
    x, w, dy, P, Q = make_conv_case(*case)
    y = torch.empty(N, P, Q, Cout, device="cuda", dtype=torch.bfloat16)
    dx = torch.zeros(N, H, W, Cin, device="cuda", dtype=torch.bfloat16)
    row = []
    dw = torch.zeros(Cout, R * R * Cin, device="cuda")
    for cl, lite in ((1, False), (2, False), (1, True)):
        pf = K.ConvPlan(K.FPROP, N, H, W, Cin, Cout, R, s, p, x, w.view(Cout, -1), y, block_n=128, cluster=cl, lite=lite)
        pd = K.ConvPlan(K.DGRAD, N, H, W, Cin, Cout, R, s, p, dx, w.view(Cout, -1), dy, addend=dx, accumulate=True, block_n=128, cluster=cl, lite=lite)
        pw = K.ConvPlan(K.WGRAD, N, H, W, Cin, Cout, R, s, p, x, None, dy, dw=dw, atomic_out=True, lite=lite)
        row.append((graph_timeit(pf.run), graph_timeit(pd.run), graph_timeit(pw.run), pf.grid))
    print(case, "fprop base/cluster2/lite %.1f / %.1f / %.1f us   dgrad %.1f / %.1f / %.1f   wgrad base/lite %.1f / %.1f  grid %s" % (
        row[0][0], row[1][0], row[2][0], row[0][1], row[1][1], row[2][1], row[0][2], row[2][2], row[0][3]))
