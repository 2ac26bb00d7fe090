"""Table of the bandwidth-class kernels from an `ncu --set full ... --page raw --csv` export (tools/ncu_elementwise.py):
  python tools/ncu_ew_summary.py gpurun_out/r02_ew_full_raw.csv [measured_GB/s] > profiles/..._summary.txt"""
import csv
import json
import sys
from pathlib import Path

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
col = {n: i for i, n in enumerate(hdr)}
peak = float(sys.argv[2]) if len(sys.argv) > 2 else None
if peak is None:
    mp = Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json"
    try:
        d = json.loads(mp.read_text())
        peak = float(d.get("hbm_gbps_burst") or d.get("hbm_gbps") or d.get("hbm_copy_gbps") or 6550.1)
    except Exception:
        peak = 6550.1


def scale(v, unit):
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "usecond": 1e-6,
                "nsecond": 1e-9, "msecond": 1e-3, "ms": 1e-3, "second": 1, "s": 1}.get(unit.lower(), 1)


def get(r, m):
    return scale(r[col[m]], units[col[m]]) if m in col and r[col[m]] not in ("", "n/a") else float("nan")


print(f"# achieved = (dram__bytes_read.sum + dram__bytes_write.sum) / gpu__time_duration.sum; frac = of the measured copy bandwidth {peak} GB/s.")
print("# ncu launches are cold-cache and serialised.")
print("kernel | grid | us | dram_read_MB | dram_write_MB | achieved_GB/s | frac_of_measured | ncu_dram_pct | regs | warps_active_pct")
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    name = r[col["Kernel Name"]].split("(")[0].replace("void ", "").replace("slider::", "")
    dur, rd, wr = get(r, "gpu__time_duration.sum"), get(r, "dram__bytes_read.sum"), get(r, "dram__bytes_write.sum")
    ach = (rd + wr) / dur * 1e-9
    print(f"{name} | {r[col['Grid Size']]} | {dur * 1e6:.1f} | {rd * 1e-6:.1f} | {wr * 1e-6:.1f} | {ach:.0f} | {ach / peak:.3f} | "
          f"{get(r, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):.1f} | {get(r, 'launch__registers_per_thread'):.0f} | "
          f"{get(r, 'sm__warps_active.avg.pct_of_peak_sustained_active'):.0f}")
