"""Summarise an `ncu --set full` report (exported with `ncu -i X.ncu-rep --page raw --csv`): one row per (kernel, grid),
the last instance of each.  Usage: ncu -i rep --page raw --csv > raw.csv; python tools/summarize_ncu_full.py raw.csv"""
import csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]


def col(*names):
    for n in names:
        for i, h in enumerate(hdr):
            if h == n:
                return i
    return None


C = {"name": col("Kernel Name"), "grid": col("Grid Size"), "t": col("gpu__time_duration.sum"),
     "rd": col("dram__bytes_read.sum"), "wr": col("dram__bytes_write.sum"),
     "dram": col("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed"),
     "tensor": col("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
                   "sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active"),
     "issue": col("smsp__issue_active.avg.pct_of_peak_sustained_active"),
     "xu": col("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
     "regs": col("launch__registers_per_thread")}
TO_US = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}
TO_MB = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}


def val(r, k, scale=1.0):
    i = C[k]
    if i is None:
        return float("nan")
    try:
        return float(r[i].replace(",", "")) * scale
    except ValueError:
        return float("nan")


last = {}
for r in rows[2:]:
    name = re.sub(r"\(.*", "", r[C["name"]]).replace("void ", "").replace("pgn::", "").replace("(anonymous namespace)::", "")
    last[(name, r[C["grid"]])] = r
print("| kernel | grid | us | dram read MB | dram write MB | dram % of peak | tensor pipe % | issue active % | MUFU pipe % | regs |")
print("|---|---|---:|---:|---:|---:|---:|---:|---:|---:|")
for (name, grid), r in last.items():
    print("| %s | %s | %.2f | %.2f | %.2f | %.1f | %.1f | %.1f | %.1f | %d |" % (
        name[:44], grid, val(r, "t", TO_US.get(units[C["t"]], 1.0)), val(r, "rd", TO_MB.get(units[C["rd"]], 1.0)),
        val(r, "wr", TO_MB.get(units[C["wr"]], 1.0)), val(r, "dram"), val(r, "tensor"), val(r, "issue"), val(r, "xu"),
        val(r, "regs")))
