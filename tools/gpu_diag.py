"""First-contact diagnostics for the tcgen05 conv kernel on a real B200.

Runs each probe in its own process (a trapped kernel kills its CUDA context) and, when an
MN-major operand gives wrong numbers, sweeps the smem-descriptor strides so one GPU call tells
us which encoding the hardware wants.  Output: gpurun_out/diag.txt
"""
import itertools
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

PROBE = r'''
import sys, json, torch
sys.path.insert(0, {root!r}); sys.path.insert(0, {root!r} + "/tests")
from gpu_util import fp32_mode, make_conv_case, ref_conv, ref_conv_grads, rel_err
from ldnn_b200.ops import native_ops as K
fp32_mode()
mode, case, ov, kw = {mode}, {case}, {ov}, {kw}
N, H, W, Cin, Cout, R, s, p = case
x, w, dy, P, Q = make_conv_case(*case)
if mode == 0:
    y = torch.zeros(N, P, Q, Cout, device="cuda", dtype=torch.bfloat16)
    K.ConvPlan(0, N, H, W, Cin, Cout, R, s, p, x, w.view(Cout, -1), y, desc_override=ov, **kw).run()
    torch.cuda.synchronize(); err = rel_err(y, ref_conv(x, w, s, p))
else:
    dxr, dwr = ref_conv_grads(x, w, dy, s, p)
    if mode == 1:
        dx = torch.zeros(N, H, W, Cin, device="cuda", dtype=torch.bfloat16)
        K.ConvPlan(1, N, H, W, Cin, Cout, R, s, p, dx, w.view(Cout, -1), dy, addend=dx, accumulate=True, desc_override=ov, **kw).run()
        torch.cuda.synchronize(); err = rel_err(dx, dxr)
    else:
        dw = torch.zeros(Cout, R * R * Cin, device="cuda", dtype=torch.float32)
        K.ConvPlan(2, N, H, W, Cin, Cout, R, s, p, x, None, dy, dw=dw, desc_override=ov, **kw).run()
        torch.cuda.synchronize(); err = rel_err(dw.view(Cout, R, R, Cin), dwr)
print("RESULT", json.dumps(err))
'''


def probe(mode, case, ov=None, **kw):
    code = PROBE.format(root=ROOT, mode=mode, case=case, ov=ov, kw=kw)
    try:
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    except subprocess.TimeoutExpired:
        return "TIMEOUT"
    for line in r.stdout.splitlines():
        if line.startswith("RESULT"):
            return float(json.loads(line.split(" ", 1)[1]))
    return "FAIL: " + (r.stderr.strip().splitlines() or ["?"])[-1][:160]


def main():
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    out = open(os.path.join(ROOT, "gpurun_out", "diag.txt"), "w")

    def log(*a):
        s = " ".join(str(x) for x in a)
        print(s, flush=True)
        out.write(s + "\n"); out.flush()

    small = (2, 8, 8, 64, 64, 1, 1, 0)
    mid = (8, 8, 8, 128, 128, 3, 1, 1)
    names = {0: "fprop", 1: "dgrad", 2: "wgrad"}
    good = {}
    for mode in (0, 1, 2):
        for case in (small, mid):
            r = probe(mode, case)
            log(names[mode], case, "default ->", r)
            good[(mode, case)] = isinstance(r, float) and r < 2e-2
    cands = [(8192, 1024), (1024, 8192), (128, 1024), (1024, 128), (8192, 128), (128, 8192), (16, 1024), (2048, 1024),
             (1024, 2048)]
    if not good[(1, mid)]:
        for lbo, sbo in cands:
            log("dgrad sweep b(lbo,sbo)=", (lbo, sbo), probe(1, mid, ov=(0, 0, lbo, sbo)))
    if not good[(2, mid)]:
        for (al, as_), (bl, bs) in itertools.product(cands[:6], cands[:6]):
            log("wgrad sweep a=", (al, as_), "b=", (bl, bs), probe(2, mid, ov=(al, as_, bl, bs)))
    out.close()


if __name__ == "__main__":
    main()
