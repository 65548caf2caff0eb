"""Throughput of the headline workload (1920x1080, default Config, device-resident inputs, every frame slot busy) and the
single-frame stage times, for the environment knobs this process was started with (PSIFT_MARCH_MIN, PSIFT_MARCH_WPS,
PSIFT_BLUR_SMALL_MAX ...).  One line of output, so that a shell loop over knob values gives a table.
usage: knob_rate.py [frames]"""
import os, sys, time, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from popsift_b200 import capi, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1500
w, h = 1920, 1080
host = [synth.gen(w, h, 1000 + i) for i in range(8)]
imgs = [torch.from_numpy(x).cuda() for x in host]
with capi.Context(capi.default_config()) as ctx:
    nslot = ctx.slot_count() if hasattr(ctx, "slot_count") else 6
    ts = []
    for i in range(10):
        ctx.extract(host[0]); ts.append(ctx.timing())
    ts = ts[3:]
    st = {k: round(statistics.median(t[k] for t in ts), 4) for k in ("pyramid", "extrema", "orientation", "descriptor") if k in ts[0]}

    def loop(m):
        q, nf = [], 0
        for i in range(m):
            q.append(ctx.submit_dev_u8(imgs[i % 8].data_ptr(), w, h, w))
            if len(q) >= nslot:
                j = q.pop(0); a, b = ctx.wait(j, copy=False); ctx.free(j); nf += a
        while q:
            j = q.pop(0); a, b = ctx.wait(j, copy=False); ctx.free(j); nf += a
        return nf
    loop(100); torch.cuda.synchronize()
    best = 0.0
    for rep in range(3):
        t0 = time.perf_counter(); loop(n); torch.cuda.synchronize(); dt = time.perf_counter() - t0
        best = max(best, n / dt)
knobs = {k: v for k, v in os.environ.items() if k.startswith("PSIFT_")}
print("%s  frames/s %.1f  stages %s" % (knobs, best, st), flush=True)
