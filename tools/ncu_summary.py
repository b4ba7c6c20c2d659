"""Key metrics per kernel from an `ncu --set full` report (run here, no GPU):
    python tools/ncu_summary.py gpurun_out/step_tf32.ncu-rep > profiles/ncu_full_key_metrics_tf32.csv
One row per distinct kernel (template arguments kept), averaged over its launches in the report."""
import collections
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__waves_per_multiprocessor"]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr = rows[0]
    col = {k: next((i for i, h in enumerate(hdr) if h == k), None) for k in KEYS}
    name_i = hdr.index("Kernel Name")
    agg = collections.OrderedDict()
    for r in rows[2:]:                      # rows[1] = units
        if len(r) != len(hdr):
            continue
        a = agg.setdefault(r[name_i], {"n": 0, **{k: 0.0 for k in KEYS}})
        a["n"] += 1
        for k, i in col.items():
            if i is None:
                continue
            try:
                a[k] += float(r[i].replace(",", ""))
            except ValueError:
                pass
    w = csv.writer(sys.stdout)
    units = rows[1]
    w.writerow(["kernel", "launches"] + [f"{k} [{units[col[k]]}]" if col[k] is not None else k for k in KEYS])
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1]["gpu__time_duration.sum"]):
        w.writerow([name, a["n"]] + [round(a[k] / a["n"], 3) for k in KEYS])


if __name__ == "__main__":
    main()
