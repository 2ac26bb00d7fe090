"""Summarises an `ncu --page raw --csv` export: per kernel name, launches, time, DRAM bytes, tensor-pipe activity.
  python tools/ncu_traffic.py gpurun_out/r01_gemm_full_raw.csv [out.json]
"""
import csv
import json
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
units = rows[1]
col = {n: i for i, n in enumerate(hdr)}
want = {"dur": "gpu__time_duration.sum", "rd": "dram__bytes_read.sum", "wr": "dram__bytes_write.sum",
        "tensor": "sm__inst_executed_pipe_tensor.sum", "tc_active": "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "l2_rd": "lts__t_bytes_equiv_l1sectormiss_pipe_lsu_mem_global_op_ld.sum", "sm_thr": "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram_thr": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"}


def scale(v, unit):
    v = float(v.replace(",", ""))
    u = unit.lower()
    mult = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "usecond": 1e-6, "nsecond": 1e-9,
            "msecond": 1e-3, "ms": 1e-3, "second": 1, "s": 1}.get(u, 1)
    return v * mult


agg = defaultdict(lambda: defaultdict(float))
per = []
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    name = r[col["Kernel Name"]].split("(")[0]
    a = agg[name]
    a["launches"] += 1
    rec = {"kernel": name, "grid": r[col["Grid Size"]] if "Grid Size" in col else None}
    for k, m in want.items():
        if m in col and r[col[m]] not in ("", "n/a"):
            v = scale(r[col[m]], units[col[m]])
            a[k] += v
            rec[k] = v
    per.append(rec)
out = {}
for name, a in agg.items():
    n = a["launches"]
    out[name] = {"launches": int(n), "sum_duration_us": round(a["dur"] * 1e6, 1),
                 "dram_read_bytes": a["rd"], "dram_write_bytes": a["wr"],
                 "avg_dram_bytes_per_launch": (a["rd"] + a["wr"]) / n,
                 "avg_sm_throughput_pct": round(a["sm_thr"] / n, 1), "avg_dram_throughput_pct": round(a["dram_thr"] / n, 1)}
print(json.dumps(out, indent=1))
if len(sys.argv) > 2:
    json.dump({"source": sys.argv[1], "kernels": out, "per_launch": per}, open(sys.argv[2], "w"))
