cd $GRAFT_REPO_ROOT
(timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | tail -5) > gpurun_out/t7_tests.log 2>&1
python tools/stage_times.py 14 > gpurun_out/t7_stage.log 2>&1
python bench.py --no-cpu --steps 12 --warmup 3 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value'],1), round(d['e2e']['value'],1), d['stage_ms'])
    else: print(l.rstrip()[-300:])
" >> gpurun_out/t7_stage.log
cat gpurun_out/t7_tests.log gpurun_out/t7_stage.log
