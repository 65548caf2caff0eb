cd $GRAFT_REPO_ROOT
for cfg in PSIFT_MARCH=0 "PSIFT_MARCH_WPS=8" "PSIFT_MARCH_WPS=16" "PSIFT_MARCH_RS=4 PSIFT_MARCH_WPS=16" "PSIFT_MARCH_WPS=8 PSIFT_MARCH_TEX=0"; do
  echo "== $cfg" >> gpurun_out/t3_bench.log
  env $cfg python bench.py --no-cpu --steps 12 --warmup 3 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value'],1), round(d['e2e']['value'],1), d['stage_ms'])
" >> gpurun_out/t3_bench.log
done
cat gpurun_out/t3_bench.log
