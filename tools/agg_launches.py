"""Aggregate an ncu `gpu__time_duration` launch list (csv) for the last training step."""
import collections, csv, sys
path = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/launches.csv"
lines = [l for l in open(path) if l.startswith('"')]
r = csv.reader(lines); hdr = next(r)
ik, iv, ig = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size")
rows = [(row[ik], float(row[iv].replace(",", "")), row[ig]) for row in r]
idx = [i for i, (k, _, _) in enumerate(rows) if "im2col_stem" in k]      # first kernel of every step
step = rows[idx[-2]: idx[-1]]
tot = sum(v for _, v, _ in step)
print(f"kernels in step {len(step)}  total {tot / 1e3:.1f} us (ncu, serialised, cold caches)")
agg = collections.defaultdict(lambda: [0, 0.0])
for k, v, g in step:
    name = k.split("(")[0][:70]
    agg[name][0] += 1; agg[name][1] += v
for k, (c, v) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{v / 1e3:9.1f} us {100 * v / tot:5.1f}% {c:4d}x  avg {v / c / 1e3:7.1f}  {k}")
