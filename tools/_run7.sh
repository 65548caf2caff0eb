cd $GRAFT_REPO_ROOT
(timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | tail -8) > gpurun_out/t7_tests.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:"k_descriptors|k_extrema" --launch-skip 3 --launch-count 3 -f -o gpurun_out/desc python tools/one_frame.py 3 > gpurun_out/t7_ncu.log 2>&1
cat gpurun_out/t7_tests.log; tail -2 gpurun_out/t7_ncu.log
