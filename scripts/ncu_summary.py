"""Summarise an `ncu --set full` capture (.ncu-rep) of bench.py: per kernel launch duration, DRAM bytes and the pipe /
memory utilisation figures the roofline discussion uses.  Writes profiles/<tag>_ncu_full_summary.{csv,json}.
    python scripts/ncu_summary.py gpurun_out/r01_full.ncu-rep r01"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, tag = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, body = rows[0], rows[1], rows[2:]
col = {n: i for i, n in enumerate(hdr)}
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size", "launch__registers_per_thread",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_tensor.sum"]
want = [w for w in want if w in col]


def num(r, name):
    v = r[col[name]].replace(",", "")
    try:
        x = float(v)
    except ValueError:
        return None
    u = units[col[name]].lower()
    scale = {"kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "byte": 1.0, "msecond": 1e3, "usecond": 1.0, "nsecond": 1e-3, "second": 1e6}
    return x * scale.get(u, 1.0)


out_csv = os.path.join(ROOT, "profiles", f"{tag}_ncu_full_summary.csv")
per = {}
with open(out_csv, "w", newline="") as f:
    w = csv.writer(f)
    w.writerow(["id", "kernel"] + want)
    for r in body:
        if len(r) <= col["Kernel Name"]:
            continue
        name = r[col["Kernel Name"]]
        vals = [num(r, n) for n in want]
        w.writerow([r[col["ID"]], name[:100]] + vals)
        key = ("chain" if "chain_kernel" in name else "gemm" if "gemm_raw_kernel" in name else "wgrad_gemm" if "gemm_kernel" in name else
               "stack" if "stack_kernel" in name else "jet_inputs" if "jet_inputs" in name else "jet_update" if "jet_update" in name else None)
        if key:
            d = dict(zip(want, vals))
            per.setdefault(key, []).append(d)
summary = {"source": f"profiles/{tag}_ncu_full_summary.csv (ncu --set full --clock-control none of `bench.py --steps 2 --warmup 3`)"}
for key, lst in per.items():
    n = len(lst)
    dram = [(d.get("dram__bytes_read.sum") or 0) + (d.get("dram__bytes_write.sum") or 0) for d in lst]
    summary[f"{key}_launches"] = n
    summary[f"{key}_dram_bytes_per_launch"] = sum(dram) / n
    summary[f"{key}_avg_us"] = sum(d["gpu__time_duration.sum"] for d in lst) / n
    for m in ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
              "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"):
        if m in lst[0] and lst[0][m] is not None:
            summary[f"{key}_{m.split('.')[0]}_pct_avg"] = sum(d[m] for d in lst) / n
json.dump(summary, open(os.path.join(ROOT, "profiles", f"{tag}_ncu_full_summary.json"), "w"), indent=1)
print(json.dumps(summary, indent=1))
