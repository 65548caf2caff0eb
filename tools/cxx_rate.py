"""Rate of the C++ classes (PopSift::enqueue / SiftJob::get) at 1080p for several producer counts / windows.
   python tools/cxx_rate.py [n_devices]     TEST INFRASTRUCTURE"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from popsift_b200 import capi, synth
nd = int(sys.argv[1]) if len(sys.argv) > 1 else 1
imgs = [synth.gen(1920, 1080, 1000 + i) for i in range(16)]
for prod, win in ((1, 8), (2, 6), (4, 4), (8, 3)):
    r = capi.cxx_stream(imgs, 100 * nd, 1500 * nd, n_devices=nd, n_producers=prod * (1 if nd == 1 else nd // 2 or 1), window=win)
    print("devices %d producers %d window %d: %.1f images/s (%.3f s, %d feat)" % (r["devices"], prod, win, r["images_per_sec"], r["seconds"], r["n_feat"]), flush=True)
