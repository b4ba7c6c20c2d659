#!/usr/bin/env python
"""SASS evidence from the built library (no GPU needed):
    python tools/sass_report.py            # writes profiles/sass/*.sass and profiles/sass_mnemonics.txt
One listing per kernel that runs in a training step or an exchange (both precisions, every tile width the tuner
uses, the CTA-pair variant), and for EVERY kernel of the library a line with the counts of the mnemonics that prove
the Blackwell paths: UTC*MMA (tcgen05.mma), UTMALDG (TMA loads), LDTM (tcgen05.ld), STRONG.SYS loads/stores (peer
memory), LDGMC/STGMC? (multimem), RED (atomics)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "ldnn_b200", "lib", "libldnn_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
KEEP = re.compile(r"conv_gemm_kernel<(float|__nv_bfloat16), [012], (64|128|256), 1, false>"      # every mode x tile width
                  r"|conv_gemm_kernel<(float|__nv_bfloat16), [01], 64, 1, true>"                      # lite launches in the step
                  r"|conv_gemm_kernel<(float|__nv_bfloat16), 0, (128|256), -2, false>"                # CTA-pair fprop
                  r"|bn_|head_|optim_kernel|cast_bf16|im2col|maxpool|allreduce|ring_|gossip|broadcast|allgather|barrier_kernel")
PAT = collections.OrderedDict([("UTCxMMA", r"\bUTC\w*MMA"), ("UTMALDG", r"\bUTMALDG"), ("LDTM", r"\bLDTM"),
                               ("UTCBAR", r"\bUTCBAR"), ("SYS ld", r"\bLDG\S*\.STRONG\.SYS|\bLD\.E\S*\.STRONG\.SYS"),
                               ("SYS st", r"\bSTG\S*\.STRONG\.SYS|\bST\.E\S*\.STRONG\.SYS"),
                               ("multimem", r"\bLDGMC|\bSTGMC|\bREDGMC|MULTIMEM"), ("RED/ATOM", r"\bRED\b|\bREDG|\bATOMG|\bATOM\b"),
                               ("MEMBAR.SYS", r"MEMBAR\S*\.SYS"), ("HMMA", r"\bHMMA")])


def short(name):
    s = re.sub(r"\((int|bool)\)", "", name.replace("void ", "").replace("ldnn::", ""))
    s = s.split("(")[0].replace("__nv_bfloat16", "bf16").replace(" ", "")
    return re.sub(r"[^A-Za-z0-9_<>,.-]", "_", s)[:100]


def main():
    dump = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    parts = re.split(r"\n\s*Function : (\S+)\n", dump)
    os.makedirs(OUT, exist_ok=True)
    for f in os.listdir(OUT):
        os.remove(os.path.join(OUT, f))
    rows = []
    for mangled, body in zip(parts[1::2], parts[2::2]):
        dem = subprocess.run(["c++filt", mangled], capture_output=True, text=True).stdout.strip()
        name = short(dem)
        counts = {k: len(re.findall(p, body)) for k, p in PAT.items()}
        nlines = sum(1 for l in body.splitlines() if re.search(r"/\*[0-9a-f]{4}\*/", l))
        rows.append((name, nlines, counts))
        if KEEP.search(dem):
            with open(os.path.join(OUT, name.replace("<", "_").replace(">", "").replace(",", "_") + ".sass"), "w") as f:
                # instruction text only (the encoding words after each instruction double the size and say nothing)
                lines = (re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l) for l in body.splitlines())
                text = "\n".join(re.sub(r"^\s+(/\*[0-9a-f]{4,}\*/)\s+", r"\1 ", l) for l in lines if l.strip())
                f.write(f"// {dem}\n" + text + "\n")
    with open(os.path.join(ROOT, "profiles", "sass_mnemonics.txt"), "w") as f:
        f.write("# cuobjdump -sass ldnn_b200/lib/libldnn_b200.so: instruction count and Blackwell-path mnemonics per kernel\n")
        f.write(f"{'kernel':<70} {'instrs':>7} " + " ".join(f"{k:>10}" for k in PAT) + "\n")
        for name, n, c in sorted(rows):
            f.write(f"{name:<70} {n:>7} " + " ".join(f"{c[k]:>10}" for k in PAT) + "\n")
    print(f"{len(rows)} kernels, {len(os.listdir(OUT))} listings in profiles/sass")


if __name__ == "__main__":
    main()
