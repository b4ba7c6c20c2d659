#!/usr/bin/env python
"""Achieved fraction of roofline for every measured hot op, from files already under profiles/ (no GPU needed):
    python tools/roofline_report.py > profiles/roofline.md
Roofline of an op = the slower of (FLOPs at the measured cuBLAS bf16 rate) and (compulsory bytes at the measured
HBM copy rate / NVLink peer-copy rate) — denominators from MEASURED_PEAKS.json (driver-written)."""
import ast
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = lambda *a: os.path.join(ROOT, *a)
LINK_GBS = 770.0            # measured peer copy per direction per GPU (profiles/exchange_sweep_8gpu.txt header run)


def main():
    try:
        peaks = json.load(open(P("MEASURED_PEAKS.json")))
    except OSError:      # driver-written file absent: the values it held when profiles/ was generated
        peaks = {"hbm_gbs": 6554.6, "bf16_tflops": 1650.7, "bf16_tflops_sustained": 1363.5}
    hbm, tf = peaks["hbm_gbs"], peaks["bf16_tflops_sustained"]
    print("# Roofline fractions (round 1)\n")
    print(f"Denominators: HBM copy {hbm:.0f} GB/s, cuBLAS bf16 sustained {tf:.0f} TFLOP/s (burst {peaks['bf16_tflops']:.0f}) "
          f"— `MEASURED_PEAKS.json`; NVLink peer copy {LINK_GBS:.0f} GB/s per direction per GPU (measured).\n")

    # ---- convolutions (profiles/microbench_kernels.txt: graph-replayed, steady state) -----------------------
    print("## Convolution kernels, repo model shapes, batch 64 (`profiles/microbench_kernels.txt`)\n")
    print("| N,H,W,Cin,Cout,R,stride,pad | GFLOP | compulsory MB (x+w+y, bf16) | roofline us | fprop us (frac) | dgrad us (frac) | wgrad us (frac) |")
    print("|---|---|---|---|---|---|---|")
    tot_t, tot_roof = 0.0, 0.0
    for line in open(P("profiles", "microbench_kernels.txt")):
        m = re.match(r"\((.*?)\)\s+([\d.]+)\s+([\d.]+)\s+([\d.]+)", line)
        if not m:
            continue
        N, H, W, Cin, Cout, R, stride, pad = ast.literal_eval("(" + m.group(1) + ")")
        Pq = (H + 2 * pad - R) // stride + 1
        flop = 2.0 * N * Pq * Pq * Cout * Cin * R * R
        byts = 2.0 * (N * H * W * Cin + Cout * Cin * R * R + N * Pq * Pq * Cout)
        roof = max(flop / (tf * 1e12), byts / (hbm * 1e9)) * 1e6
        ts = [float(m.group(i)) for i in (2, 3, 4)]
        tot_t += sum(ts)
        tot_roof += 3 * roof
        cells = " | ".join(f"{t:.1f} ({roof / t:.2f})" for t in ts)
        print(f"| {N},{H},{W},{Cin},{Cout},{R},{stride},{pad} | {flop / 1e9:.2f} | {byts / 1e6:.1f} | {roof:.1f} | {cells} |")
    print(f"\nSum over these shapes: {tot_t:.0f} us measured vs {tot_roof:.0f} us roofline -> {tot_roof / tot_t:.2f}.  These layers are "
          "launch/fill-latency bound at batch 64 (a whole layer is 4-16 us); see DESIGN.md 'Known gaps' and NEXT.md.\n")

    # ---- optimizer -----------------------------------------------------------------------------------------------
    for line in open(P("profiles", "microbench_kernels.txt")):
        m = re.match(r"adam over (\d+) elems: ([\d.]+) us\s+\(([\d.]+) TB/s\)", line)
        if m:
            print("## Flat Adam (fp32 p/g/m/v read, p/m/v + bf16 shadow written)\n")
            print(f"{int(m.group(1)) / 1e6:.1f} M elements in {m.group(2)} us = {m.group(3)} TB/s -> "
                  f"{float(m.group(3)) * 1e3 / hbm:.2f} of the HBM copy roofline.\n")

    # ---- step ----------------------------------------------------------------------------------------------------
    b = json.load(open(P("profiles", "bench_1gpu_final.json")))
    print("## Whole step, 1 GPU (`profiles/bench_1gpu_final.json`)\n")
    ms = b["ms_per_step"]
    # fwd FLOPs of the repo model at 32x32: counted from the spec (conv + fc), x3 for fwd+dgrad+wgrad
    import sys
    sys.path.insert(0, ROOT)
    from ldnn_b200.models import enhanced_cnn_spec
    fwd = float(enhanced_cnn_spec().forward_flops(1))
    if fwd:
        tot = 3 * fwd * b["config"]["per_gpu_batch"]
        print(f"{tot / 1e9:.0f} GFLOP per step (3 x forward) in {ms:.3f} ms = {tot / ms / 1e9:.0f} TFLOP/s -> "
              f"{tot / ms / 1e9 / tf:.2f} of the cuBLAS bf16 rate; e2e {b['e2e']['value']:.0f} vs device-timed {b['value']:.0f} samples/s.\n")
    else:
        print(f"{ms:.3f} ms per step; e2e {b['e2e']['value']:.0f} vs device-timed {b['value']:.0f} samples/s.\n")

    # ---- exchange ------------------------------------------------------------------------------------------------
    print("## Fused exchange kernels, 8 GPUs, fp32 payload (`profiles/exchange_sweep_8gpu.txt`)\n")
    print("Bus bandwidth = 2(W-1)/W x bytes / time; roofline = bytes over NVLink at the measured peer-copy rate "
          "(compute is negligible: one FMA per 4 bytes).\n")
    print("| payload | two-shot GB/s (frac) | ring GB/s (frac) | double ring GB/s (frac) | NCCL + scale GB/s (frac) |")
    print("|---|---|---|---|---|")
    for line in open(P("profiles", "exchange_sweep_8gpu.txt")):
        m = re.match(r"\s*(\d+)\s+([\d.]+)\|\s*(\d+)\s+([\d.]+)\|\s*(\d+)\s+([\d.]+)\|\s*(\d+)\s+([\d.]+)\|\s*(\d+)", line)
        if not m:
            continue
        nbytes = int(m.group(1))
        if nbytes < (1 << 20) or (nbytes & (nbytes - 1)) or nbytes not in (1 << 20, 1 << 24, 1 << 26, 1 << 27, 1 << 28, 1 << 30):
            continue
        bw = [float(m.group(i)) for i in (2, 4, 6, 8)]
        label = f"{nbytes >> 20} MB"
        print(f"| {label} | " + " | ".join(f"{v:.0f} ({v / LINK_GBS:.2f})" for v in bw) + " |")
    print("\nThe repo model's gradient is 178 MB: between the 128 MB and 256 MB rows.")


if __name__ == "__main__":
    main()
