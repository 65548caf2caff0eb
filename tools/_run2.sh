cd $GRAFT_REPO_ROOT
PSIFT_MARCH_WPS=16 ncu --set full --import-source on --clock-control none -k regex:k_blur_march --launch-skip 12 --launch-count 6 -f -o gpurun_out/march python tools/one_frame.py 3 > gpurun_out/t2_ncu.log 2>&1
tail -3 gpurun_out/t2_ncu.log
