"""Matcher timing on a GPU box: psift_match (host call: launch + D2H of the result) against the reference's
compute_distance kernel (oracle/_ref/ref_driver --match-desc, kernel + sync only) on the same descriptor sets.
TEST/MEASUREMENT TOOL.   gpurun -- python tools/match_bench.py [n ...]"""
import json, os, subprocess, sys, tempfile, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from popsift_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRV = os.path.join(ROOT, "oracle", "_ref", "ref_driver")

def main():
    sizes = [int(a) for a in sys.argv[1:]] or [3000, 12000]
    rng = np.random.default_rng(0)
    for n in sizes:
        v = rng.gamma(0.6, 1.0, (2 * n, 128)).astype(np.float32)
        v = np.sqrt(v / v.sum(1, keepdims=True)).astype(np.float32)
        l, r = v[:n], v[n:]
        with capi.DevFeatures.from_host(l) as a, capi.DevFeatures.from_host(r) as b:
            a.match(b)
            t0 = time.perf_counter()
            for _ in range(5): m, d = a.match(b)
            ms = (time.perf_counter() - t0) / 5 * 1e3
        out = {"n": n, "psift_match_ms": round(ms, 3), "pairs_per_s": round(n * n / ms * 1e3)}
        if os.path.exists(DRV):
            with tempfile.TemporaryDirectory() as td:
                lf, rf, of = (os.path.join(td, x) for x in ("l.f32", "r.f32", "m.bin"))
                l.tofile(lf); r.tofile(rf)
                p = subprocess.run([DRV, "--match-desc", lf, rf, "--out", of], capture_output=True, text=True, timeout=600)
                ref = json.loads(p.stdout.strip().splitlines()[-1])
                out["reference_kernel_ms"] = ref["ms"]
                out["same_result"] = bool(np.array_equal(np.fromfile(of, "<i4").reshape(-1, 3), m))
        print(json.dumps(out), flush=True)

if __name__ == "__main__":
    main()
