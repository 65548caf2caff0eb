#!/bin/bash
# Launch list (durations only) of the third 1080p frame.  usage (under gpurun): bash tools/ncu_list.sh TAG [launches per frame]
T=$1; N=${2:-57}
PSIFT_GRAPH=0 python tools/one_frame.py 3 > gpurun_out/of.log 2>&1 &&
PSIFT_GRAPH=0 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip $((2*N)) -c $N --csv --log-file gpurun_out/${T}_list.csv python tools/one_frame.py 3 > gpurun_out/ncu_l.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/${T}_list.csv")) if len(r)>10]
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value"); gi=hdr.index("Grid Size")
tot=0
for r in rows[1:]:
    v=float(r[vi].replace(",",""))/1000.0; tot+=v
    if not r[ki].startswith("k_blur2"): print("%-40s %-16s %8.2f us"%(r[ki][:40], r[gi], v))
print("sum %.1f us over %d launches"%(tot,len(rows)-1))
PY
