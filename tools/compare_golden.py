"""Run the new build over the golden cases on a GPU and print the parity report against the
committed reference outputs (tests/golden/ref).  TEST INFRASTRUCTURE.
    gpurun -- python tools/compare_golden.py [case ...]
"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import refio, parity
from popsift_b200 import capi
from tools import golden_cases as gc

only = sys.argv[1:]
out = {}
for case, (img, flags, planes) in gc.CASES.items():
    if only and case not in only: continue
    fb = os.path.join(refio.GOLD, "ref", case + ".feat.bin")
    if not os.path.exists(fb): continue
    w, h, _ = gc.IMAGES[img]
    im = np.fromfile(gc.image_path(img), np.uint8).reshape(h, w)
    try:
        cfg = capi.config_from_flags(flags)
        with capi.Context(cfg) as ctx:
            mine = ctx.extract(im)
            ref = refio.RefFeatures(fb)
            rep = parity.compare(mine, ref, norm_multi=cfg.norm_multi)
            if planes:
                rep["planes"] = parity.compare_planes(ctx, refio.RefPlanes(os.path.join(refio.GOLD, "ref", case + ".planes.bin")))
    except Exception as e:
        rep = {"error": repr(e)}
    out[case] = rep
    print(case, json.dumps(rep), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "compare_golden.json"), "w"), indent=1)
