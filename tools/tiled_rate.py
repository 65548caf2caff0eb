"""One giant image (BASELINE config 5: 16384 x 16384, --downsampling 0): un-split on one GPU against psift_extract_banded
over all visible GPUs (one context per GPU), equality by order-independent digests.  TEST INFRASTRUCTURE.
   python tools/tiled_rate.py [size] [n_ctx] [reps]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from popsift_b200 import capi, synth, banded
size = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
ndev = capi.lib().psift_device_count()
n_ctx = int(sys.argv[2]) if len(sys.argv) > 2 else ndev
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
tile = synth.gen(2048, 2048, 5)
img = np.tile(tile, (size // 2048, size // 2048))
cfg = capi.default_config(downsampling=0); cfg.max_extrema = 400000
t = []
with capi.Context(cfg) as ctx:
    whole = ctx.extract(img)
    for _ in range(reps):
        t0 = time.perf_counter(); j = ctx.submit(img); ctx.wait(j, copy=False); ctx.free(j); t.append(time.perf_counter() - t0)
    tm = ctx.timing()
print("un-split, 1 GPU: %s s; stages ms %s; %d keypoints %d descriptors" % (["%.3f" % x for x in t], {k: round(v, 2) for k, v in tm.items()}, whole.n_feat, whole.n_desc), flush=True)
dw = banded.digest(whole)
for ex in (1, 0):
    os.environ["PSIFT_BAND_EXCHANGE"] = str(ex)
    ctxs = [capi.Context(cfg, device=i % ndev) for i in range(n_ctx)]
    tb = []
    got = capi.extract_banded(ctxs, img)
    for _ in range(reps):
        t0 = time.perf_counter(); capi.extract_banded(ctxs, img, copy=False); tb.append(time.perf_counter() - t0)
    for c in ctxs: c.close()
    print("banded, %d contexts on %d GPUs, halo_exchange %d, split %d: %s s; equal to un-split: %s" % (n_ctx, ndev, got.halo_exchange, got.split, ["%.3f" % x for x in tb],
          (got.n_feat, got.n_desc) == (whole.n_feat, whole.n_desc) and banded.digest(got) == dw), flush=True)
