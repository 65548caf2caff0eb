#!/bin/bash
# One GPU box: the round's evidence set into gpurun_out/.  usage (under gpurun): bash tools/evidence_run.sh TAG bench|frame
T=${1:-r02}
if [ "${2:-bench}" = bench ]; then
  # the bench line, the reference arm, and the ncu launch list of the (short) bench command
  python bench.py > gpurun_out/${T}_bench_1gpu.json 2> gpurun_out/${T}_bench_1gpu.err
  python bench.py --impl reference > gpurun_out/${T}_bench_reference_arm.json 2> gpurun_out/${T}_ref.err
  python bench.py --steps 2 --warmup 3 --no-cpu --no-tiled > gpurun_out/plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/${T}_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-tiled > gpurun_out/ncu_b.log 2>&1
  python - <<PY
import json
d=json.load(open("gpurun_out/${T}_bench_1gpu.json")); print(d["value"], d["e2e"]["value"], d["stage_ms"], d["roofline"]["frac"], d["cpu_baseline"]["value"], d["gpu_launches"], d["tiled16k"]["ms_per_image"] if d.get("tiled16k") else None)
r=json.load(open("gpurun_out/${T}_bench_reference_arm.json")); print(r["value"], r["runs_images_per_sec"], r["host_cpu"])
PY
else
  # every launch of one warm frame with all metrics; only the raw-page CSV comes back (the report itself is > 64 MiB)
  PSIFT_GRAPH=0 python tools/one_frame.py 3 > gpurun_out/of.log 2>&1 &&
  PSIFT_GRAPH=0 ncu --set full --clock-control none --launch-skip 112 -c 56 -o /tmp/${T}_frame python tools/one_frame.py 3 > gpurun_out/ncu_f.log 2>&1
  ncu -i /tmp/${T}_frame.ncu-rep --page raw --csv > gpurun_out/${T}_frame_raw.csv 2>/dev/null
  tail -2 gpurun_out/ncu_f.log; ls -la /tmp/${T}_frame.ncu-rep gpurun_out/${T}_frame_raw.csv
fi
