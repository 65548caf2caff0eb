cd $GRAFT_REPO_ROOT
python tools/run_ref_golden.py g64_scaledirect g256_scaledirect g64_scaledirect_rel g256_relative > gpurun_out/t5_gold.log 2>&1
(timeout 600 python -m pytest tests/test_parity_gpu.py -x -q -k "relative or scaledirect" 2>&1 | tail -25) > gpurun_out/t5_rel.log 2>&1
python tools/submit_cost.py > gpurun_out/t5_submit.log 2>&1
cat gpurun_out/t5_gold.log gpurun_out/t5_rel.log gpurun_out/t5_submit.log
