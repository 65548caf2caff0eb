cd $GRAFT_REPO_ROOT
(PSIFT_MARCH_MIN=0 PSIFT_MARCH_TEX=1 timeout 900 python -m pytest tests/test_parity_gpu.py -x -q -k "golden or banded or edge" 2>&1 | tail -5) > gpurun_out/t4_tests.log 2>&1
(timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | tail -5) >> gpurun_out/t4_tests.log 2>&1
python bench.py > gpurun_out/r01d_bench_1gpu.json 2> gpurun_out/t4_bench.err
python bench.py --impl reference --steps 4 --warmup 1 > gpurun_out/r01d_bench_reference_arm.json 2>> gpurun_out/t4_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r01d_launches_bench.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/t4_ncu.log 2>&1
cat gpurun_out/t4_tests.log; tail -3 gpurun_out/t4_bench.err; cat gpurun_out/r01d_bench_1gpu.json gpurun_out/r01d_bench_reference_arm.json
