cd $GRAFT_REPO_ROOT
(PSIFT_MARCH_MIN=0 PSIFT_MARCH_TEX=1 timeout 900 python -m pytest tests/test_parity_gpu.py -x -q -k "golden or banded or edge" 2>&1 | tail -15) > gpurun_out/t1_march_all.log 2>&1
(PSIFT_MARCH_MIN=0 PSIFT_MARCH_TEX=1 PSIFT_MARCH_RS=4 timeout 900 python -m pytest tests/test_parity_gpu.py -x -q -k "golden or banded or edge" 2>&1 | tail -15) >> gpurun_out/t1_march_all.log 2>&1
for cfg in PSIFT_MARCH=0 "PSIFT_MARCH_WPS=8" "PSIFT_MARCH_WPS=12" "PSIFT_MARCH_WPS=16" "PSIFT_MARCH_RS=4 PSIFT_MARCH_WPS=12" "PSIFT_MARCH_RS=4 PSIFT_MARCH_WPS=16"  PSIFT_MARCH_TEX=0 PSIFT_MARCH_MIN=1500000; do
  echo "== $cfg" >> gpurun_out/t1_stage.log
  env $cfg python tools/stage_times.py 14 >> gpurun_out/t1_stage.log 2>&1
done
PSIFT_MARCH_WPS=8 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/t1_launch_march.csv python tools/one_frame.py 3 > gpurun_out/t1_ncu1.log 2>&1
PSIFT_MARCH_WPS=16 PSIFT_MARCH_RS=4 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/t1_launch_march16.csv python tools/one_frame.py 3 > gpurun_out/t1_ncu2.log 2>&1
PSIFT_MARCH_WPS=8 ncu --set full --import-source on --clock-control none -k regex:k_blur_march --launch-skip 12 --launch-count 6 -f -o gpurun_out/march8 python tools/one_frame.py 3 > gpurun_out/t2_ncu.log 2>&1
cat gpurun_out/t1_march_all.log gpurun_out/t1_stage.log
