"""Median per-stage times (CUDA events on the frame's stream) for single frames.  usage: stage_times.py [n]"""
import os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from popsift_b200 import capi, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 12
img = synth.gen(1920, 1080, 1000)
with capi.Context(capi.default_config()) as ctx:
    ts = []
    for i in range(n):
        ctx.extract(img); ts.append(ctx.timing())
    ts = ts[3:]
    print({k: round(statistics.median(t[k] for t in ts), 4) for k in ts[0] if k not in ("upload", "download", "total")})
