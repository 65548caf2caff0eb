"""Frames per second of BASELINE.json configs 1-3 (device-resident inputs, all frame slots busy), next to the headline config 4.
usage: config_rates.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from popsift_b200 import capi, synth
NF = int(os.environ.get("PSIFT_SLOTS", "6"))
CONFIGS = [("c1 640x480 default", 640, 480, 1, {}),
           ("c2 1920x1080 vlfeat igrid", 1920, 1080, 2, dict(mode="vlfeat", gauss_mode="vlfeat", downsampling=-1, desc_mode="igrid")),
           ("c3 3840x2160 opencv rootsift filter 4000", 3840, 2160, 3, dict(mode="opencv", gauss_mode="opencv", norm_mode="RootSift", filter_max_extrema=4000, filter_grid_size=4, filter_sort="down")),
           ("c3b 3840x2160 opencv rootsift no filter", 3840, 2160, 3, dict(mode="opencv", gauss_mode="opencv", norm_mode="RootSift")),
           ("c4 1920x1080 default", 1920, 1080, 1000, {})]
for name, w, h, seed, kw in CONFIGS:
    imgs = [torch.from_numpy(synth.gen(w, h, seed + i)).cuda() for i in range(4)]
    with capi.Context(capi.default_config(**kw)) as ctx:
        def loop(n):
            q, nf, nd = [], 0, 0
            for i in range(n):
                q.append(ctx.submit_dev_u8(imgs[i % 4].data_ptr(), w, h, w))
                if len(q) >= NF:
                    j = q.pop(0); a, b = ctx.wait(j, copy=False); ctx.free(j); nf += a; nd += b
            while q:
                j = q.pop(0); a, b = ctx.wait(j, copy=False); ctx.free(j); nf += a; nd += b
            return nf, nd
        n = 200 if w * h < 4e6 else 80
        loop(30); torch.cuda.synchronize()
        t0 = time.perf_counter(); nf, nd = loop(n); torch.cuda.synchronize(); dt = time.perf_counter() - t0
        print("%-44s %8.1f frames/s  %7.3f ms/frame  %7.0f keypoints %7.0f descriptors per frame" % (name, n / dt, 1e3 * dt / n, nf / n, nd / n), flush=True)
