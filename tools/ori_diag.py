"""Diagnostic: num_ori differences between this build and the reference on one seeded image; the reference is run
several times to show its own run-to-run spread (float atomics in its orientation histogram).  TEST INFRASTRUCTURE.
   python tools/ori_diag.py W H SEED [ref flags...]"""
import os, sys, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import refio
from popsift_b200 import capi, synth
from tools import golden_cases as gc
w, h, seed = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]); flags = sys.argv[4:]
img = synth.gen(w, h, seed); img.tofile("/tmp/od.u8")
refs = []
for k in range(4):
    subprocess.run([os.path.join(ROOT, "oracle/_ref/ref_driver"), "--in", "/tmp/od.u8", "--w", str(w), "--h", str(h), "--out", "/tmp/od%d" % k] + flags, check=True, capture_output=True)
    refs.append(refio.RefFeatures("/tmp/od%d.feat.bin" % k))
with capi.Context(capi.config_from_flags(gc.config_flags(flags))) as ctx:
    mine = ctx.extract(img)
om = refio.canonical_order(mine.ext); me = mine.ext[om]
print("n_feat mine", mine.n_feat, "refs", [r.n_feat for r in refs], "n_desc mine", mine.n_desc, "refs", [r.n_desc for r in refs])
res = [r.ext[refio.canonical_order(r.ext)] for r in refs]
for i in range(len(me)):
    ns = [int(r["num_ori"][i]) for r in res]
    if len(set(ns + [int(me["num_ori"][i])])) > 1:
        print("kp", i, "oct", me["octave"][i], "x,y", me["x"][i], me["y"][i], "sigma", me["sigma"][i], "mine", int(me["num_ori"][i]), me["ori"][i], "refs", ns, [r["ori"][i] for r in res])
