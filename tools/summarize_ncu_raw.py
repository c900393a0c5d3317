"""Table from an `ncu -i rep --page raw --csv` export: per captured launch the kernel, its duration, DRAM bytes
and the derived GB/s against MEASURED_PEAKS.json, tensor-pipe utilisation, registers, issue-slot utilisation."""
import csv, json, os, re, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rows = list(csv.reader(l for l in open(sys.argv[1]) if not l.startswith("==")))
h, units = rows[0], rows[1]
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
sc = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}
def val(r, k):
    if k not in h: return None
    i = h.index(k)
    try: return float(r[i].replace(",", "")) * sc.get(units[i], 1.0)
    except ValueError: return None
print(f"{'kernel':44s} {'grid':>8s} {'us':>8s} {'dram MB':>8s} {'GB/s':>7s} {'of peak':>7s} {'tensor%':>8s} {'issue%':>7s} {'regs':>5s}")
for r in rows[2:]:
    name = re.sub(r"\(.*", "", r[h.index("Kernel Name")]).replace("void ", "").replace("mt::", "")
    t = val(r, "gpu__time_duration.sum")
    b = (val(r, "dram__bytes_read.sum") or 0) + (val(r, "dram__bytes_write.sum") or 0)
    tp = val(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active")
    isu = val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active")
    regs = val(r, "launch__registers_per_thread")
    grid = r[h.index("Grid Size")] if "Grid Size" in h else ""
    print(f"{name[:44]:44s} {grid:>8s} {t*1e6:8.1f} {b/1e6:8.1f} {b/t/1e9:7.0f} {b/t/1e9/peak:7.2f} "
          f"{(tp if tp is not None else 0):8.1f} {(isu if isu is not None else 0):7.1f} {int(regs or 0):5d}")
