#!/usr/bin/env python
"""Achieved fraction of roofline for every measured hot op, from files already under profiles/ (no GPU needed):
    python tools/roofline_report.py > profiles/roofline.md
Every input file carries its own clock record.  Denominators: MEASURED_PEAKS.json (driver-written: HBM copy, cuBLAS
bf16 sustained; TF32 = half of the bf16 rate, the hardware ratio) and the measured 770 GB/s NVLink peer copy."""
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = lambda *a: os.path.join(ROOT, *a)
LINK = 770.0


def main():
    try:
        peaks = json.load(open(P("MEASURED_PEAKS.json")))
    except OSError:
        peaks = {"hbm_gbs": 6554.6, "bf16_tflops": 1650.7, "bf16_tflops_sustained": 1363.5}
    hbm, tf = peaks["hbm_gbs"], peaks["bf16_tflops_sustained"]
    print("# Roofline fractions (round 2)\n")
    print(f"Denominators (of measured): HBM copy {hbm:.0f} GB/s, cuBLAS bf16 sustained {tf:.0f} TFLOP/s (burst "
          f"{peaks['bf16_tflops']:.0f}), TF32 taken as half of it; NVLink peer copy {LINK:.0f} GB/s per direction per GPU.\n")

    # ---- kernels (profiles/microbench_kernels_r2.txt carries the per-shape table + its clock record) ------------
    mb = open(P("profiles", "microbench_kernels_r2.txt")).read()
    print("## Convolution / BN / Adam kernels, repo model shapes, batch 64 — `profiles/microbench_kernels_r2.txt`\n")
    for dt in ("tf32", "bf16"):
        m = re.search(rf"== conv {dt}.*?\n(.*?)\nsum (\d+) us vs roofline (\d+) us -> ([\d.]+)", mb, re.S)
        if not m:
            continue
        print(f"**conv {dt}** (warm L2, graph-replayed; fprop / dgrad / wgrad us, then fraction of roofline per pass)\n")
        print("| N,H,W,Cin,Cout,R,stride,pad | fprop | dgrad | wgrad | roofline us | frac f / d / w |")
        print("|---|---|---|---|---|---|")
        for line in m.group(1).splitlines():
            mm = re.match(r"\((.*?)\)\s+([\d.]+)\s+([\d.]+)\s+([\d.]+) \|\s+([\d.]+) \| ([\d.]+) ([\d.]+) ([\d.]+)", line)
            if mm:
                g = mm.groups()
                print(f"| {g[0]} | {g[1]} | {g[2]} | {g[3]} | {g[4]} | {g[5]} / {g[6]} / {g[7]} |")
        print(f"\nSum over these shapes: {m.group(2)} us measured vs {m.group(3)} us roofline -> **{m.group(4)}**.\n")
    for line in mb.splitlines():
        if line.startswith("flat Adam"):
            print("* " + line)
    clk = re.search(r"# clocks during these measurements: (.*)", mb)
    print(f"\nClock record of that file: `{clk.group(1) if clk else 'missing'}`.\n")

    # ---- whole step ------------------------------------------------------------------------------------------------
    print("## Whole training step (`profiles/bench_r2.jsonl`; 207 GFLOP per 64-sample step = 3 x forward)\n")
    print("| arm | GPUs | topology / schedule | samples/s | ms/step | e2e samples/s | fraction of the tensor roofline | clocks |")
    print("|---|---|---|---|---|---|---|---|")
    for l in open(P("profiles", "bench_r2.jsonl")):
        d = json.loads(l)
        c = d["config"]
        if c["model"] != "enhanced_cnn" or c.get("shard_ratios"):
            continue
        rate = tf * (0.5 if d["dtype"] == "tf32" else 1.0) * 1e12
        frac = 207e9 / (d["ms_per_step"] * 1e-3) / rate
        ck = d.get("clocks") or {}
        print(f"| {d['impl']} ({d['dtype']}) | {d['n_gpus']} | {c['topology']} / {c['sync']} | {d['value']:.0f} | "
              f"{d['ms_per_step']:.3f} | {d['e2e']['value']:.0f} | {frac:.2f} | {ck.get('sm_mhz')} MHz, {ck.get('reasons')} |")
        for k, v in ((d.get("extra") or {}).get("runs") or {}).items():
            if "value" in v and not v.get("headline"):
                print(f"| {d['impl']} ({d['dtype']}), extra | {d['n_gpus']} | {k.replace('/', ' / ')}{' (restated)' if v.get('restated') else ''} | "
                      f"{v['value']:.0f} | {v['ms_per_step']:.3f} | | | |")
        b = (d.get("extra") or {}).get("bf16")
        if b and "value" in b:
            print(f"| ours (bf16), extra | {d['n_gpus']} | {c['topology']} / {c['sync']} | {b['value']:.0f} | {b['ms_per_step']:.3f} | "
                  f"{b['e2e']['value']:.0f} | {207e9 / (b['ms_per_step'] * 1e-3) / (tf * 1e12):.2f} | |")
    print()

    # ---- fused exchanges ---------------------------------------------------------------------------------------------
    print("## Fused exchange kernels (`profiles/exchange_sweep_{2,4,8}_r2.json`, fp32 payload)\n")
    print("Roofline of an exchange = the slower of (bytes each way over NVLink, 2(W-1)/W x payload, at 770 GB/s) and (HBM "
          "time of the local passes); `update` = the fused reduce + scale + Adam + parameter push, against NCCL all_reduce "
          "+ scale + torch fused Adam.  Cells: bus bandwidth GB/s (fraction of roofline).\n")
    for n in (2, 4, 8):
        try:
            blob = json.load(open(P("profiles", f"exchange_sweep_{n}_r2.json")))
        except OSError:
            continue
        rows = blob["rows"]
        impls = [t for t in ("allreduce", "allreduce_nvls", "ring", "double_ring", "nccl") if any(r["impl"] == t for r in rows)]
        print(f"**{n} GPUs** — clocks during the sweep: `{json.dumps(blob['clocks'])}`\n")
        for mode in ("reduce", "update"):
            print(f"| {mode}: payload | " + " | ".join({"allreduce": "two-shot P2P", "allreduce_nvls": "NVLS (in-switch)",
                                                       "ring": "ring", "double_ring": "double ring", "nccl": "NCCL arm"}[t]
                                                      for t in impls) + " |")
            print("|---|" + "---|" * len(impls))
            for mbs in (1, 4, 16, 64, 128, 256, 512):
                cells = []
                for t in impls:
                    r = [x for x in rows if x["impl"] == t and x["mode"] == mode and x["bytes"] == mbs << 20]
                    cells.append(f"{r[0]['busbw_GBs']:.0f} ({r[0]['roofline_frac']:.2f})" if r else "-")
                if any(c != "-" for c in cells):
                    print(f"| {mbs} MB | " + " | ".join(cells) + " |")
            print()
    print("The repo model's gradient is 178 MB (between the 128 MB and 256 MB rows); the per-step schedule exchanges it in "
          "32 MB (all-reduce) or 64 MB (rings) buckets.")


if __name__ == "__main__":
    main()
