"""Markdown table from an `ncu --page raw --csv` export: one row per kernel (first captured launch).
usage: python scripts/summarise_ncu_raw.py raw.csv > table.md"""
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}


def col(r, name, default=""):
    i = ix.get(name)
    return r[i] if i is not None and i < len(r) else default


seen = set()
print("| kernel | grid x block | time (us) | DRAM read+write (MB) | dram throughput % of peak | regs/thread | dyn smem (KB) | fp64 pipe % |")
print("|---|---|---|---|---|---|---|---|")
for r in rows[2:]:
    name = col(r, "Kernel Name")
    short = re.sub(r"\(.*", "", name).replace("void ", "").replace("<unnamed>::", "")
    if not short or short in seen:
        continue
    seen.add(short)

    def f(metric, scale=1.0, default=float("nan")):
        try:
            return float(col(r, metric).replace(",", "")) * scale
        except ValueError:
            return default
    t = f("gpu__time_duration.sum")
    tu = rows[1][ix["gpu__time_duration.sum"]] if "gpu__time_duration.sum" in ix else "us"
    t = t * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}.get(tu, 1.0)

    def bytes_mb(metric):
        v = f(metric)
        u = rows[1][ix[metric]] if metric in ix else "byte"
        return v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1e-6)
    mb = bytes_mb("dram__bytes_read.sum") + bytes_mb("dram__bytes_write.sum")
    smem = float("nan")
    if "launch__shared_mem_per_block_dynamic" in ix:
        su = rows[1][ix["launch__shared_mem_per_block_dynamic"]]
        smem = f("launch__shared_mem_per_block_dynamic") * {"byte/block": 1e-3, "Kbyte/block": 1.0, "Mbyte/block": 1e3}.get(su, 1e-3)
    print(f"| `{short}` | {col(r, 'launch__grid_size')} x {col(r, 'launch__block_size')} | {t:.1f} | {mb:.1f} | "
          f"{f('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):.1f} | {col(r, 'launch__registers_per_thread')} | "
          f"{smem:.1f} | {f('sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active'):.1f} |")
