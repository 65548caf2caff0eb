"""How many orientations of the golden cases are bit-equal to the reference's (PSIFT_ORI_ORDER=0/1).  TEST INFRASTRUCTURE."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, refio, parity
from popsift_b200 import capi
from tools import golden_cases as gc
tot = [0, 0, 0, 0]
for case in ["g256_default", "g640_default", "g256_opencv", "g256_down0", "g256_levels4", "g64_default", "check_default", "g256_l2"]:
    img, flags, _ = gc.CASES[case]
    w, h, _ = gc.IMAGES[img]
    im = np.fromfile(gc.image_path(img), np.uint8).reshape(h, w)
    with capi.Context(capi.config_from_flags(gc.config_flags(flags))) as ctx:
        mine = ctx.extract(im)
    rep = parity.compare(mine, refio.RefFeatures(os.path.join(refio.GOLD, "ref", case + ".feat.bin")))
    print(case, "ok_int", rep["ok_int"], "ori_bitexact", rep.get("ori_bitexact"), "desc_bitexact", rep.get("desc_bitexact"), "max_dtheta", rep.get("max_dtheta"))
    if "ori_bitexact" in rep:
        tot[0] += rep["ori_bitexact"][0]; tot[1] += rep["ori_bitexact"][1]; tot[2] += rep["desc_bitexact"][0]; tot[3] += rep["desc_bitexact"][1]
print("TOTAL order=%s ori %d/%d desc %d/%d" % (os.environ.get("PSIFT_ORI_ORDER", "0"), *tot))
