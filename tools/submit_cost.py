"""Host time spent inside psift_submit_* per frame (is the submitting thread the bottleneck?).  usage: submit_cost.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from popsift_b200 import capi, synth
imgs = [synth.gen(1920, 1080, 1000 + i) for i in range(4)]
dev = [torch.from_numpy(im).cuda() for im in imgs]
with capi.Context(capi.default_config()) as ctx:
    for name, sub in (("dev", lambda i: ctx.submit_dev_u8(dev[i % 4].data_ptr(), 1920, 1080, 1920)), ("host", lambda i: ctx.submit(imgs[i % 4]))):
        q = []; ts = 0.0; tw = 0.0; n = 200
        t0 = time.perf_counter()
        for i in range(n):
            a = time.perf_counter(); q.append(sub(i)); ts += time.perf_counter() - a
            if len(q) >= 4:
                a = time.perf_counter(); j = q.pop(0); ctx.wait(j, copy=False); ctx.free(j); tw += time.perf_counter() - a
        while q:
            j = q.pop(0); ctx.wait(j, copy=False); ctx.free(j)
        tot = time.perf_counter() - t0
        print("%s: total %.3f ms/frame, in submit %.3f ms/frame, in wait %.3f ms/frame" % (name, 1e3 * tot / n, 1e3 * ts / n, 1e3 * tw / n))
