#!/bin/bash
# One kernel of a warm 1080p frame under ncu --set full with source: raw page and SASS/source page as CSV into gpurun_out/.
# usage (under gpurun): bash tools/ncu_kernel.sh TAG KERNEL_REGEX [launch-skip]
T=$1; K=$2; S=${3:-2}
PSIFT_GRAPH=0 python tools/one_frame.py 3 > gpurun_out/of.log 2>&1 &&
PSIFT_GRAPH=0 ncu --set full --import-source on --clock-control none -k "regex:$K" --launch-skip $S -c 1 -o /tmp/$T python tools/one_frame.py 3 > gpurun_out/ncu_k.log 2>&1
ncu -i /tmp/$T.ncu-rep --page raw --csv > gpurun_out/${T}_raw.csv 2>/dev/null
ncu -i /tmp/$T.ncu-rep --page source --csv > gpurun_out/${T}_source.csv 2>/dev/null
tail -2 gpurun_out/ncu_k.log; ls -la gpurun_out/${T}_raw.csv gpurun_out/${T}_source.csv
