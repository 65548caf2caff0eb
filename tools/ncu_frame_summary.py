"""Summarise an `ncu --set full` report of one frame (ncu -i REPORT --page raw --csv > raw.csv) as one line per launch.
usage: python tools/ncu_frame_summary.py raw.csv"""
import csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, data = rows[0], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
def g(r, k, d=0.0):
    try: return float(r[ix[k]].replace(",", ""))
    except Exception: return d
cols = [("gpu__time_duration.sum", "time us", 1e-3 if rows[1][ix["gpu__time_duration.sum"]] in ("ns", "nsecond") else 1.0),
        ("dram__bytes_read.sum", "dram rd MB", None), ("dram__bytes_write.sum", "dram wr MB", None),
        ("smsp__inst_issued.sum", "warp-instr M", 1e-6), ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps act %", 1),
        ("smsp__issue_active.avg.pct", "issue %", 1), ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma pipe %", 1),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu %", 1),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex %", 1), ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts %", 1),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram %", 1), ("launch__registers_per_thread", "regs", 1)]
def mb(r, k):
    u = rows[1][ix[k]]
    return g(r, k) * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1e-6)
print("# kernel | grid | " + " | ".join(c[1] for c in cols))
tot = 0.0
for r in data:
    name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("void ", "").strip()
    vals = []
    for k, label, sc in cols:
        if k not in ix: vals.append("   n/a"); continue
        v = mb(r, k) if sc is None else g(r, k) * sc
        vals.append("%7.2f" % v)
    tot += g(r, cols[0][0]) * cols[0][2]
    print("%-34s | %6s | %s" % (name[:34], r[ix["launch__grid_size"]], " | ".join(vals)))
print("# sum of launch durations: %.1f us" % tot)
