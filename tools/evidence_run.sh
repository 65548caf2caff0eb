#!/bin/bash
# One GPU box: the round's evidence set into gpurun_out/ (bench line, reference arm, ncu launch list of the bench command,
# full-metric capture of every launch of one frame).  usage (under gpurun): bash tools/evidence_run.sh TAG
T=${1:-r02}
python bench.py > gpurun_out/${T}_bench_1gpu.json 2> gpurun_out/${T}_bench_1gpu.err
python bench.py --impl reference > gpurun_out/${T}_bench_reference_arm.json 2> gpurun_out/${T}_ref.err
python bench.py --steps 2 --warmup 3 --no-cpu --no-tiled > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/${T}_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-tiled > gpurun_out/ncu_b.log 2>&1
PSIFT_GRAPH=0 python tools/one_frame.py 3 > gpurun_out/of.log 2>&1 &&
PSIFT_GRAPH=0 ncu --set full --clock-control none --launch-skip 116 -c 58 -o gpurun_out/${T}_frame python tools/one_frame.py 3 > gpurun_out/ncu_f.log 2>&1
tail -2 gpurun_out/ncu_f.log
python - <<PY
import json
d=json.load(open("gpurun_out/${T}_bench_1gpu.json")); print(d["value"], d["e2e"]["value"], d["stage_ms"], d["roofline"]["frac"], d["cpu_baseline"]["value"], d["gpu_launches"], d["tiled16k"]["ms_per_image"] if d.get("tiled16k") else None)
r=json.load(open("gpurun_out/${T}_bench_reference_arm.json")); print(r["value"], r["runs_images_per_sec"], r["host_cpu"])
PY
