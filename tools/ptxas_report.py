#!/usr/bin/env python
"""Registers / stack / spills of every kernel, from `nvcc -Xptxas -v` (no GPU needed).
    python tools/ptxas_report.py > profiles/ptxas_resources.txt"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    print("# nvcc -gencode arch=compute_100a,code=sm_100a -O3 -Xptxas -v; one line per kernel")
    for src in ("conv", "bn", "head_optim", "comm"):
        cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-Xptxas", "-v",
               "-I", os.path.join(ROOT, "include"), "-c", os.path.join(ROOT, "csrc", src + ".cu"), "-o", os.devnull]
        out = subprocess.run(cmd, capture_output=True, text=True).stderr
        name, st = None, ("?", "?", "?")
        for line in out.splitlines():
            m = re.search(r"Compiling entry function '([^']*)'", line)
            if m:
                name = m.group(1)
                continue
            m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
            if m:
                st = m.groups()
                continue
            m = re.search(r"Used (\d+) registers", line)
            if m and name:
                short = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.split("(")[0].strip()
                sm = re.search(r"(\d+) bytes smem", line)
                print(f"{src + '.cu':14s} {short[:58]:58s} regs={m.group(1):>3s} stack={st[0]:>4s} spill_st={st[1]:>3s} "
                      f"spill_ld={st[2]:>3s} static_smem={sm.group(1) if sm else '0':>5s}")
                name = None


if __name__ == "__main__":
    sys.exit(main())
