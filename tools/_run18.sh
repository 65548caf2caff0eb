cd $GRAFT_REPO_ROOT
python -m pytest tests/test_parity_gpu.py -x -q -k "device_resident" 2>&1 | tail -3
python bench.py --no-cpu --steps 12 --warmup 3 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value'],1), d['e2e'])
"
