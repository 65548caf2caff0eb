"""Run the reference CUDA build (oracle/_ref/ref_driver) over tools/golden_cases.py on a GPU
box and leave the outputs in gpurun_out/golden/.  TEST INFRASTRUCTURE.

    gpurun -- python tools/run_ref_golden.py
"""
import os, subprocess, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools import golden_cases as gc

ROOT = os.path.dirname(gc.HERE)
OUT = os.path.join(ROOT, "gpurun_out", "golden")
os.makedirs(OUT, exist_ok=True)
drv = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
log = {}
only = sys.argv[1:]
for case, (img, flags, planes) in gc.CASES.items():
    if only and case not in only:
        continue
    w, h, _ = gc.IMAGES[img]
    src = gc.image_path(img)
    if gc.is_float(flags):
        src = os.path.join(OUT, "_" + img + ".f32")
        gc.float_image(img).tofile(src)
    cmd = [drv, "--in", src, "--w", str(w), "--h", str(h), "--out", os.path.join(OUT, case)] + flags
    if planes:
        cmd += ["--dump-planes", "--dump-tables"]
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=20)
    except subprocess.TimeoutExpired:
        print(case, "TIMEOUT", flush=True); log[case] = {"rc": -9}; continue
    log[case] = {"rc": r.returncode, "stdout": r.stdout.strip(), "stderr": r.stderr.strip()[-400:]}
    print(case, r.returncode, r.stdout.strip(), r.stderr.strip()[-200:], flush=True)
# matcher goldens: descriptor sets come from the committed extraction goldens
sys.path.insert(0, os.path.join(ROOT, "tests"))
import refio
for case, (lc, rc_) in gc.MATCH_CASES.items():
    if only and case not in only:
        continue
    files = []
    for c in (lc, rc_):
        fn = os.path.join(OUT, "_desc_" + c + ".f32")
        refio.RefFeatures(os.path.join(refio.GOLD, "ref", c + ".feat.bin")).desc.astype("<f4").tofile(fn)
        files.append(fn)
    r = subprocess.run([drv, "--match-desc", files[0], files[1], "--out", os.path.join(OUT, case + ".match.bin")], capture_output=True, text=True, timeout=60)
    log[case] = {"rc": r.returncode, "stdout": r.stdout.strip(), "stderr": r.stderr.strip()[-400:]}
    print(case, r.returncode, r.stdout.strip(), r.stderr.strip()[-200:], flush=True)
json.dump(log, open(os.path.join(OUT, "log.json"), "w"), indent=1)
