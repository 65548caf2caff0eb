"""Host-side cost of the submit/wait loop: frames/s of a tiny image (the GPU is idle most of the time, so this is the rate
the calling thread can sustain) and of 1080p frames, host and device inputs.  usage: submit_cost.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from popsift_b200 import capi, synth
NF = int(os.environ.get("PSIFT_SLOTS", "6"))
def loop(ctx, sub, n):
    q = []
    for i in range(n):
        q.append(sub(i))
        if len(q) >= NF:
            j = q.pop(0); ctx.wait(j, copy=False); ctx.free(j)
    while q:
        j = q.pop(0); ctx.wait(j, copy=False); ctx.free(j)
for w, h in ((64, 48), (1920, 1080)):
    imgs = [synth.gen(w, h, 1000 + i) for i in range(4)]
    dev = [torch.from_numpy(im).cuda() for im in imgs]
    with capi.Context(capi.default_config()) as ctx:
        for name, sub in (("dev", lambda i: ctx.submit_dev_u8(dev[i % 4].data_ptr(), w, h, w)), ("host", lambda i: ctx.submit(imgs[i % 4]))):
            loop(ctx, sub, 60)
            t0 = time.perf_counter(); loop(ctx, sub, 400); dt = time.perf_counter() - t0
            print("%dx%d %s: %.3f ms/frame (%.0f frames/s)" % (w, h, name, 1e3 * dt / 400, 400 / dt))
