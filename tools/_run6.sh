cd $GRAFT_REPO_ROOT
for cfg in PSIFT_SLOTS=4 PSIFT_SLOTS=6; do
  echo "== $cfg" >> gpurun_out/t6_bench.log
  env $cfg python bench.py --no-cpu --steps 12 --warmup 3 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value'],1), round(d['e2e']['value'],1), d['stage_ms'], d['gpu_launches'])
    else: print(l.rstrip()[-300:])
" >> gpurun_out/t6_bench.log
done
python tools/submit_cost.py >> gpurun_out/t6_bench.log 2>&1
cat gpurun_out/t6_bench.log
