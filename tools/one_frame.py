"""Run a few 1080p frames through the context (profiling target for ncu).  usage: one_frame.py [n] [desc_mode]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from popsift_b200 import capi, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
kw = {"desc_mode": sys.argv[2]} if len(sys.argv) > 2 else {}
img = synth.gen(1920, 1080, 1000)
with capi.Context(capi.default_config(**kw)) as ctx:
    for i in range(n):
        r = ctx.extract(img)
    print(r.n_feat, r.n_desc, ctx.timing())
