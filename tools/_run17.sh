cd $GRAFT_REPO_ROOT
for cfg in PSIFT_BLUR_SMALL_MAX=3000000 PSIFT_BLUR_SMALL_MAX=1500000 PSIFT_BLUR_SMALL_MAX=400000; do
echo "== $cfg" >> gpurun_out/t17.log
env $cfg python bench.py --no-cpu --steps 12 --warmup 3 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value'],1), round(d['e2e']['value'],1), d['stage_ms']['pyramid'])
" >> gpurun_out/t17.log
done
cat gpurun_out/t17.log
