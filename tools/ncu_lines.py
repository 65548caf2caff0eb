"""Per-source-line instruction and stall-sample shares of one kernel: joins the SASS page of an ncu report
(ncu -i X.ncu-rep --page source --csv) with the line table of the cubin (nvdisasm -g -c).  TEST INFRASTRUCTURE.
   python tools/ncu_lines.py sass.csv disasm.txt kernel_substring [min_share_percent]"""
import csv, re, sys
sass_csv, dis, kern = sys.argv[1], sys.argv[2], sys.argv[3]
thr = float(sys.argv[4]) if len(sys.argv) > 4 else 0.5
rows = list(csv.reader(open(sass_csv)))
hdr = rows[1]
ci, cs, ct = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Avg. Threads Executed")
ins = [(r[1].strip(), int(r[ci]), int(r[cs]), float(r[ct] or 0)) for r in rows[2:] if len(r) > ci and r[ci].isdigit()]
# line table: walk the function's text section in order
lines, cur, on = [], None, False
for l in open(dis):
    if l.startswith(".text."):
        on = kern in l
        continue
    if not on: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l): lines.append(cur)
print("sass rows", len(ins), "disasm instructions", len(lines))
n = min(len(ins), len(lines))
agg = {}
tot_i = sum(x[1] for x in ins); tot_s = sum(x[2] for x in ins)
for k in range(n):
    a = agg.setdefault(lines[k], [0, 0, 0.0])
    a[0] += ins[k][1]; a[1] += ins[k][2]; a[2] += ins[k][1] * ins[k][3]
print("total warp instructions %d, samples %d" % (tot_i, tot_s))
for key in sorted(agg, key=lambda k: (k is None, k)):
    a = agg[key]
    if 100.0 * a[0] / tot_i >= thr or 100.0 * a[1] / max(tot_s, 1) >= thr:
        print("%-22s instr %6.2f%%  samples %6.2f%%  avg threads %5.1f" % ("%s:%d" % key if key else "?", 100.0 * a[0] / tot_i, 100.0 * a[1] / max(tot_s, 1), a[2] / max(a[0], 1)))
