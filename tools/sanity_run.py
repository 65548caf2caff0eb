"""Small run through every code path added in round 2 (for compute-sanitizer): default + igrid + grid-filter frames, the
linear-time candidate sort with a crowded octave, halo-exchange banded extraction (loop and igrid), the C++ classes
through psift_cxx_stream, psift_wait_take blocks, the no-allocation matcher.  TEST INFRASTRUCTURE."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from popsift_b200 import capi, synth
img = synth.gen(640, 480, 1)
for kw in ({}, {"desc_mode": "igrid", "mode": "vlfeat"}, {"filter_max_extrema": 200, "filter_grid_size": 3, "filter_sort": "down"}, {"threshold": 0.005}):
    with capi.Context(capi.default_config(**kw)) as ctx:
        a = ctx.extract(img); b = ctx.extract(img)
        assert a.n_feat == b.n_feat and a.desc.tobytes() == b.desc.tobytes()
        print(kw, a.n_feat, a.n_desc, flush=True)
tall = synth.gen(512, 2048, 9)
for flags in ([ "--downsampling", "0"], ["--downsampling", "0", "--desc-mode", "igrid"]):
    cfg = capi.config_from_flags(flags)
    with capi.Context(cfg) as ctx: whole = ctx.extract(tall)
    ctxs = [capi.Context(cfg) for _ in range(3)]
    got = capi.extract_banded(ctxs, tall, max_seed_pixels=128 * 512)
    for c in ctxs: c.close()
    assert (got.n_feat, got.n_desc) == (whole.n_feat, whole.n_desc), (got.n_feat, whole.n_feat)
    print("banded", flags, got.n_feat, "exchange", got.halo_exchange, flush=True)
r = capi.cxx_stream([img, synth.gen(640, 480, 2)], 4, 12, n_devices=1, n_producers=2, window=3)
print("cxx", r["n_feat"], r["n_desc"], flush=True)
import ctypes as C
d = np.ascontiguousarray(a.desc[:300])
with capi.DevFeatures.from_host(d) as f:
    m = np.zeros((f.n_desc, 3), np.int32)
    assert capi.lib().psift_match_dev_noalloc(0, f.info.desc, f.n_desc, f.info.desc, f.n_desc, m.ctypes.data, None) == 0
    assert (m[:, 0] == np.arange(f.n_desc)).sum() > 0.9 * f.n_desc
print("sanity ok")
