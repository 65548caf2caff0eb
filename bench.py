#!/usr/bin/env python
"""bench.py -- throughput of the SIFT extraction hot path on B200 ("Measurement").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one pass of the hot path over one batch of BATCH synthetic 1920x1080 frames per GPU
(default popsift::Config: 2x upscale, 9 octaves, loop descriptors, RootSift).  Prints ONE JSON
line on rank 0.  Workload = BASELINE.json configs[3]/[1] (1080p frames through enqueue()).

  value     images/s over all GPUs, inputs already resident in HBM (psift_submit_dev_u8), results
            downloaded to pinned host memory (the download is part of the path).
  e2e       the same with HOST buffers: page-locked input through psift_submit_pinned_u8 (H2D inside);
            e2e.value_private_copy: pageable input through psift_submit_u8 (private pinned copy + H2D inside).
  roofline  pyramid stage (all k_blur2 launches of a frame): algorithmic bytes
            A_pyr = w*h + 4*L*sum(octave pixels) over the CUDA-event time of the stage.
  cpu_baseline  the CPU restatement (oracle/) timed on one host core on ONE frame -- the
            reference has no CPU SIFT path at all, so this is a port, not the reference.
  --impl reference  times the reference's own CUDA build (oracle/_ref/ref_driver) on the same GPU
            and frames; the reference is single-GPU, so only rank 0 runs it.
"""
import argparse, json, os, subprocess, sys, threading, time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H = 1920, 1080
BATCH = 25            # frames per step per GPU
N_DISTINCT = 16       # distinct synthetic frames cycled through (seeds 1000..1015)
IN_FLIGHT = int(os.environ.get("PSIFT_SLOTS", "6"))   # frames in flight = frame slots of the context
BLUR_DRAM_BYTES = 33.6e6       # dram read+write of one k_blur_march launch (3840x2160 level), ncu --set full: profiles/r02_ncu_blur_march.txt
PROFILE_SRC = "profiles/r02_ncu_blur_march.txt, profiles/r02q_ncu_frame_summary.txt"
N_PYR_LAUNCHES = 46
PYR_DRAM_BYTES = 227.4e6       # dram read+write of the 46 pyramid launches of one frame (profiles/r02q_ncu_frame_summary.txt)
REF_REPEATS = 3
CXX_PRODUCERS = 2     # caller threads per GPU that enqueue() and get() in the C++-class leg
CXX_WINDOW = 6        # jobs each of them keeps outstanding
ONE_OBJECT_PRODUCERS = 4   # caller threads per GPU when ONE PopSift object serves all GPUs of the box
REF_MAX_FRAMES = 2000


def host_cpu():
    try:
        for l in open("/proc/cpuinfo"):
            if l.startswith("model name"): return l.split(":", 1)[1].strip() + " x%d" % (os.cpu_count() or 0)
    except Exception:
        pass
    return "unknown"


def pyramid_bytes(cfg_levels=3, up=1):
    import math
    w0, h0 = W << up, H << up
    n_oct = max(int(math.floor(math.log2(min(W, H))) - 3 + (1 << up)), 1)
    tot, w, h = 0, w0, h0
    for _ in range(n_oct):
        tot += w * h
        w, h = (w + 1) // 2, (h + 1) // 2
    return W * H + 4 * (cfg_levels + 3) * tot, tot, n_oct


class ClockSampler:
    """SM clock and throttle reasons during the timed region (B200_PROFILING.md's clocks line), read
    through NVML in-process.  Spawning nvidia-smi five times a second instead takes driver locks
    often enough to slow the reference arm, which synchronises the device many times per frame,
    by 3x (measured) -- that would flatter the comparison, so it is only the fallback, once a second."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    BITS = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}

    def __init__(self, index):
        self.index, self.samples, self.stop = index, [], False
        self.t = threading.Thread(target=self.run, daemon=True)
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def sample(self):
        if self.nvml:
            n = self.nvml
            sm = n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)
            mx = n.nvmlDeviceGetMaxClockInfo(self.h, n.NVML_CLOCK_SM)
            try:
                r = n.nvmlDeviceGetCurrentClocksEventReasons(self.h)
            except Exception:
                r = n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            return [str(sm), str(mx)] + ["Active" if r & self.BITS[k] else "Not Active" for k in ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")]
        o = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                           capture_output=True, text=True, timeout=5).stdout.strip()
        return [x.strip() for x in o.split(",")] if o else None

    def run(self):
        while not self.stop:
            try:
                v = self.sample()
                if v: self.samples.append(v)
            except Exception:
                pass
            time.sleep(0.1 if self.nvml else 1.0)

    def __enter__(self): self.t.start(); return self
    def __exit__(self, *a): self.stop = True; self.t.join(timeout=6)

    def summary(self):
        sm = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        mx = [int(s[1]) for s in self.samples if s[1].isdigit()]
        reasons = set()
        for s in self.samples:
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], s[2:6]):
                if v.lower().startswith("active"): reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.samples), "source": "nvml" if self.nvml else "nvidia-smi"}


def workload_config(n_oct=9):
    """the `config` object of BOTH arms (the driver compares them key by key)"""
    return {"workload": "stream of 1920x1080 u8 frames, default popsift::Config (2x upscale -> 3840x2160 octave 0, %d octaves, 6 Gaussian levels, loop descriptors, RootSift)" % n_oct,
            "frames_per_step": BATCH, "distinct_frames": N_DISTINCT,
            "l2": "no explicit flush: every frame writes its own Gaussian pyramid (265 MB here, 752 MB in the reference), larger than the 126 MB L2"}


def frames():
    from popsift_b200 import synth
    return [synth.gen(W, H, 1000 + i) for i in range(N_DISTINCT)]


def run_stream(ctx, submit, imgs, n):
    """n frames through the context with IN_FLIGHT jobs in flight; returns (features, descriptors)."""
    q, nf, nd = [], 0, 0
    for i in range(n):
        q.append(submit(imgs[i % len(imgs)]))
        if len(q) >= IN_FLIGHT:
            j = q.pop(0); a, b = ctx.wait(j, copy=False); ctx.free(j); nf += a; nd += b
    while q:
        j = q.pop(0); a, b = ctx.wait(j, copy=False); ctx.free(j); nf += a; nd += b
    return nf, nd


def cpu_baseline():
    exe = os.path.join(ROOT, "oracle", "cpu_sift")
    if not os.path.exists(exe):
        return {"value": None, "unit": "images/s", "cores": 1, "kind": "port", "sample": "oracle/cpu_sift not built"}
    from popsift_b200 import synth
    p = "/tmp/bench_cpu_frame.u8"
    synth.gen(W, H, 1000).tofile(p)
    t0 = time.time()
    r = subprocess.run([exe, "--in", p, "--w", str(W), "--h", str(H), "--quiet"], capture_output=True, text=True)
    dt = time.time() - t0
    ok = r.returncode == 0
    return {"value": (1.0 / dt) if ok else None, "unit": "images/s", "cores": 1, "kind": "port",
            "sample": "1 frame 1920x1080 (seed 1000), default Config, single-thread C++ restatement of the reference kernels (the reference has no CPU path); %.1f s" % dt}


TILED = 16384          # BASELINE.json configs[4]: one 16384 x 16384 image, --downsampling 0


def giant_image():
    """16384 x 16384 synthetic image: an 8 x 8 mosaic of one 2048 x 2048 synth.gen() tile (generating 268 Mpx of
    independent content takes minutes; equality with the un-split run does not depend on the content)"""
    import numpy as np
    from popsift_b200 import synth
    return np.tile(synth.gen(2048, 2048, 5), (TILED // 2048, TILED // 2048))


def tiled_leg(n_gpus, steps=8, warmup=2):
    """BASELINE config 5 through psift_extract_banded: one context per GPU of THIS process, row bands with per-level halo
    exchange over NVLink, host image in, merged host results out; wall clock per image, next to the un-split run on
    one GPU; equality of the two results (order-independent digests) is asserted."""
    from popsift_b200 import capi, banded
    img = giant_image()
    cfg = capi.default_config(downsampling=0); cfg.max_extrema = 400000
    with capi.Context(cfg, device=0) as ctx:
        whole = ctx.extract(img)
        for _ in range(warmup):
            j = ctx.submit(img); ctx.wait(j, copy=False); ctx.free(j)
        t0 = time.perf_counter()
        for _ in range(steps):
            j = ctx.submit(img); ctx.wait(j, copy=False); ctx.free(j)
        t_whole = (time.perf_counter() - t0) / steps
        stage = ctx.timing()
    want = (whole.n_feat, whole.n_desc) + banded.digest(whole)
    n_ctx = max(n_gpus, 2)                    # on one GPU: two bands on the same device (exercises the exchange path)
    ctxs = [capi.Context(cfg, device=i % n_gpus) for i in range(n_ctx)]
    try:
        got = capi.extract_banded(ctxs, img)
        same = (got.n_feat, got.n_desc) + banded.digest(got) == want
        for _ in range(warmup): capi.extract_banded(ctxs, img, copy=False)
        t0 = time.perf_counter()
        for _ in range(steps): capi.extract_banded(ctxs, img, copy=False)
        t_band = (time.perf_counter() - t0) / steps
    finally:
        for c in ctxs: c.close()
    assert same, "banded result differs from the un-split run"
    return {"workload": "one 16384x16384 u8 image, default Config with --downsampling 0 (12 octaves), host image in, host results out",
            "n_gpus": n_gpus, "bands": n_ctx, "halo_exchange": got.halo_exchange, "split_octave": got.split,
            "ms_per_image": 1e3 * t_band, "images_per_sec": 1.0 / t_band,
            "unsplit_1gpu_ms_per_image": 1e3 * t_whole, "speedup_vs_unsplit": t_whole / t_band,
            "unsplit_stage_ms": {k: round(v, 3) for k, v in stage.items()},
            "keypoints": whole.n_feat, "descriptors": whole.n_desc, "equal_to_unsplit": bool(same), "steps": steps, "timing": "wall clock",
            "host": {"cpus_allowed": len(os.sched_getaffinity(0)), "loadavg": os.getloadavg()[0]}}


def bench_tiled(args, rank, world):
    """--workload tiled16k: config 5 alone.  One process drives all N GPUs (the bands of one image cooperate), so under
    torchrun rank 0 runs and the other ranks exit."""
    if rank != 0:
        return
    t = tiled_leg(args.gpus, steps=max(args.steps, 1), warmup=max(args.warmup, 1))
    line = {"metric": "images_per_sec_16384_tiled", "value": t["images_per_sec"], "unit": "images/s", "n_gpus": args.gpus, "steps": t["steps"],
            "warmup": max(args.warmup, 1), "ms_per_step": t["ms_per_image"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": {"workload": t["workload"], "bands": t["bands"]},
            "e2e": {"value": t["images_per_sec"], "unit": "images/s", "h2d_bytes_per_step": TILED * TILED,
                    "d2h_bytes_per_step": t["keypoints"] * 96 + t["descriptors"] * 512,
                    "note": "the tiled entry point takes a host image and returns host results: value and e2e are the same host-to-host figure"},
            "tiled": t}
    print(json.dumps(line))


def bench_reference(args, rank):
    if rank != 0:
        return
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
    base = {"impl": "reference", "metric": "images_per_sec_1080p", "unit": "images/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic"}
    if not os.path.exists(exe):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/ref_driver not built (run __graft_entry__.build() where /root/reference exists)"}))
        return
    imgs = frames()
    paths = []
    for i, im in enumerate(imgs):
        p = "/tmp/bench_ref_%d.u8" % i; im.tofile(p); paths += ["--in", p]
    # the reference synchronises the device many times per frame, so its rate depends on the host CPU and on what else
    # the box is doing: the whole K-step run is repeated REF_REPEATS times and the best is reported (all are listed)
    n_iters = min(args.steps * BATCH, REF_MAX_FRAMES)       # bounded sample: the arm ends within a few minutes on any box
    runs = []
    with ClockSampler(0) as cs:
        for _ in range(REF_REPEATS):
            r = subprocess.run([exe] + paths + ["--w", str(W), "--h", str(H), "--iters", str(n_iters), "--warmup", str(min(args.warmup, 4) * BATCH)],
                               capture_output=True, text=True, timeout=1500)
            if r.returncode != 0:
                print(json.dumps({"impl": "reference", "unavailable": "ref_driver failed: " + r.stderr.strip()[-200:]}))
                return
            runs.append(json.loads(r.stdout.strip().splitlines()[-1]))
    d = max(runs, key=lambda x: x["images_per_sec"])
    v = d["images_per_sec"]
    base.update({"value": v, "ms_per_step": d["ms_per_frame"] * BATCH, "ms_per_frame": d["ms_per_frame"],
                 "config": workload_config(),
                 "note": "the reference's own CUDA kernels built for sm_100 (two sm_70+ hazards fixed in a build-tree copy, oracle/Makefile.ref), driven like its worker thread: load -> step1 -> step2 -> get_descriptors, one image at a time; single GPU only",
                 "runs_images_per_sec": [x["images_per_sec"] for x in runs], "best_of": REF_REPEATS, "host_cpu": host_cpu(),
                 "features_per_frame": d["features_total"] / max(d["iters"], 1),
                 "cpu_baseline": {"value": v, "unit": "images/s", "cores": 1, "kind": "reference",
                                  "sample": "%d frames 1920x1080 (best of %d runs); the reference has NO CPU implementation of this path, so this arm is its own CUDA build on the same B200, driven by one host thread" % (n_iters, REF_REPEATS)},
                 "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "gpu_launches": None, "clocks": cs.summary()})
    print(json.dumps(base))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=240)      # 240 x 25 frames: a timed region of a little over 2 s per leg
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--workload", default="stream1080p", help="stream1080p (default, BASELINE configs[3]) or tiled16k (configs[4])")
    ap.add_argument("--no-tiled", action="store_true", help="skip the giant-image leg of the default workload")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        bench_reference(args, rank)
        return
    if args.workload == "tiled16k":
        bench_tiled(args, rank, world)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    from popsift_b200 import capi
    from popsift_b200 import dist as pdist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the extraction path has no CPU fallback")
    imgs = frames()

    # The two legs in which ONE process drives every GPU of the box run first, while the other ranks wait on a CPU-side
    # (gloo) barrier and have not created a CUDA context yet: a GPU shared with another process's live context moved
    # rank 0's result downloads at 1.3 GB/s on the 8-GPU box (143 ms for a 16384^2 frame's results instead of 2 ms).
    if world > 1:
        dist.init_process_group("gloo")                      # CPU only; the NCCL group for the per-rank legs follows
    one_object, tiled = None, None
    if rank == 0:
        if world > 1:
            # ONE PopSift object over all N GPUs of the box (its in-process dispatcher: one worker per GPU, shared queue)
            n_prod = max(2, min(ONE_OBJECT_PRODUCERS * world, (os.cpu_count() or 8) - world))   # caller threads; each worker needs a core too
            r1 = capi.cxx_stream(imgs, args.warmup * BATCH * world, args.steps * BATCH * world, first_device=0, n_devices=world,
                                 n_producers=n_prod, window=CXX_WINDOW)
            one_object = {"value": r1["images_per_sec"], "unit": "images/s", "devices": r1["devices"], "producer_threads": n_prod,
                          "what": "one PopSift object dispatching to all GPUs of this process (rank 0, before the other ranks start), wall clock"}
        if not args.no_tiled:
            try:
                tiled = tiled_leg(world)                     # BASELINE config 5: one giant image over the GPUs of the box
            except Exception as e:                           # the headline line must not depend on this leg
                tiled = {"error": repr(e)[:300]}
    if world > 1:
        dist.barrier()                                       # gloo: the waiting ranks sleep on a socket

    torch.cuda.set_device(local)
    if world > 1:
        pdist.use_group(dist.new_group(backend="nccl", device_id=torch.device("cuda", local)))   # barrier + max-over-ranks of the timed legs

    barrier = pdist.barrier

    dev_imgs = [torch.from_numpy(im).cuda() for im in imgs]
    ctx = capi.Context(capi.default_config(), device=local)
    sub_dev = lambda t: ctx.submit_dev_u8(t.data_ptr(), W, H, W)
    sub_host = lambda im: ctx.submit(im)
    pin_imgs = [torch.from_numpy(im).pin_memory() for im in imgs]          # the same frames in page-locked host memory
    sub_pin = lambda t: ctx.submit_pinned_u8(t.data_ptr(), W, H, W)

    def timed(submit, src):
        """K steps: device time between two CUDA-event marks that join every stream of the context
        (psift_debug_mark), max over ranks; the wall clock around the same region is kept as a check."""
        for _ in range(args.warmup): run_stream(ctx, submit, src, BATCH)
        torch.cuda.synchronize(); barrier(); torch.cuda.synchronize()
        ctx.mark(0)
        t0 = time.perf_counter()
        nf = nd = 0
        for _ in range(args.steps):
            a, b = run_stream(ctx, submit, src, BATCH); nf += a; nd += b
        ctx.mark(1)
        dt = ctx.mark_elapsed_ms() * 1e-3
        torch.cuda.synchronize(); barrier(); torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        dt = pdist.max_over_ranks(dt); wall = pdist.max_over_ranks(wall)
        return dt, pdist.sum_over_ranks(nf) / world, pdist.sum_over_ranks(nd) / world, wall

    l0 = ctx.launch_count()
    with ClockSampler(local) as cs:
        dt_dev, nf, nd, wall_dev = timed(sub_dev, dev_imgs)
    launches = (ctx.launch_count() - l0) * args.steps // (args.steps + args.warmup)
    b0 = ctx.d2h_bytes()
    dt_e2e, nf2, nd2, wall_e2e = timed(sub_pin, pin_imgs)
    d2h_per_step = (ctx.d2h_bytes() - b0) // (args.steps + args.warmup)
    dt_copy, _, _, _ = timed(sub_host, imgs)          # psift_submit_u8: enqueue()'s semantics, pageable input + private copy
    n_frames = args.steps * BATCH

    # END TO END through the reference's own API: the C++ classes, PopSift::enqueue(w, h, pageable pointer) / SiftJob::get()
    # (psift_cxx_stream, host/bench_shim.cpp): one PopSift object on this rank's GPU, CXX_PRODUCERS caller threads, frames
    # enqueued at the rate they retire.  Wall clock inside the library call (the classes expose no device events): it
    # contains the job copy, upload, kernels, result download and the walk over the returned features.  Max over ranks.
    torch.cuda.synchronize(); barrier()
    cxx = capi.cxx_stream(imgs, args.warmup * BATCH, n_frames, first_device=local, n_devices=1, n_producers=CXX_PRODUCERS, window=CXX_WINDOW)
    barrier()
    dt_cxx = pdist.max_over_ranks(cxx["seconds"])
    cxx_feat = pdist.sum_over_ranks(cxx["n_feat"]) / world; cxx_desc = pdist.sum_over_ranks(cxx["n_desc"]) / world
    # kernel-stage times (CUDA events on the frame's stream), one frame at a time, warm
    stage = {}
    for i in range(8):
        j = sub_dev(dev_imgs[i % len(dev_imgs)]); ctx.wait(j, copy=False); ctx.free(j)
        t = ctx.timing()
        for k, v in t.items(): stage.setdefault(k, []).append(v)
    stage = {k: float(np.median(v)) for k, v in stage.items()}
    a_pyr, sum_px, n_oct = pyramid_bytes()
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    peak = peaks.get("hbm_gbs", 6650.0)
    stage_achieved = a_pyr / (stage["pyramid"] * 1e-3) / 1e9

    # dominant kernel, per launch: the blur launches of octave 0 (levels 1..5 = k_blur_march<H,8,0>; level 0 samples
    # the input through the texture unit and is listed separately), each bracketed by CUDA events on the stream it is
    # launched on (psift_debug_time_levels).  The stream is kept full (IN_FLIGHT frames) like the timed region above.
    ctx.time_levels(True)
    lv = []
    q = []
    for i in range(24):
        q.append(sub_dev(dev_imgs[i % len(dev_imgs)]))
        if len(q) >= IN_FLIGHT:
            j = q.pop(0); ctx.wait(j, copy=False); ctx.free(j); lv.append(ctx.level_timing())
    while q:
        j = q.pop(0); ctx.wait(j, copy=False); ctx.free(j); lv.append(ctx.level_timing())
    lv_alone = []
    for i in range(8):
        j = sub_dev(dev_imgs[i % len(dev_imgs)]); ctx.wait(j, copy=False); ctx.free(j); lv_alone.append(ctx.level_timing())
    ctx.time_levels(False)
    lv = np.median(np.array(lv[4:]), axis=0); lv_alone = np.median(np.array(lv_alone[2:]), axis=0)
    w0, h0 = W * 2, H * 2
    lvl_bytes = [W * H + 4 * w0 * h0] + [8 * w0 * h0] * 5          # level 0: u8 input + one fp32 plane; levels 1..5: read one plane, write one (A_pyr2 model)
    lvl_bytes[3] += 4 * (w0 // 2) * (h0 // 2)                          # level 3 also writes the next octave's seed
    dom_bytes = float(np.mean(lvl_bytes[1:])); dom_ms = float(np.mean(lv_alone[1:])); dom_ms_busy = float(np.mean(lv[1:]))
    achieved = dom_bytes / (dom_ms * 1e-3) / 1e9

    # fp32 pipe: a bit-exact separable filter with H one-sided taps costs H adds + (H+1) FMAs (horizontal, symmetric taps
    # folded) + (2H+1) FMAs (vertical) = 4H+2 lane-instructions per pixel; peak = SMs x 128 lanes x SM clock
    taps = [5, 7, 8, 10, 13]
    n_sm = torch.cuda.get_device_properties(local).multi_processor_count
    fp32_peak = n_sm * 128 * peaks.get("sm_max_mhz", 1965.0) * 1e6
    fp32_ops = [(4 * t + 2) * w0 * h0 for t in taps]
    fp32_frac = float(sum(fp32_ops) / (sum(lv_alone[1:]) * 1e-3) / fp32_peak)
    strict_bytes = 4.0 * w0 * h0
    roofline = {"bound": "hbm", "achieved": stage_achieved, "peak": peak, "unit": "GB/s", "frac": stage_achieved / peak,
                "what": "pyramid stage, SURVEY 8d: A_pyr = w*h + 4*L*sum(px) (input read once, every Gaussian level written once) over the stage's duration on the frame's stream (CUDA events, one warm frame at a time, median of 8)",
                "algorithmic_bytes": a_pyr, "ms": stage["pyramid"], "launches": N_PYR_LAUNCHES,
                "traffic": PYR_DRAM_BYTES, "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum summed over the pyramid launches of one frame, ncu --set full, " + PROFILE_SRC,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback",
                "dominant_kernel": {
                    "kernel": "k_blur_march<H,8,0> (one Gaussian level of octave 0, 3840x2160: the 5 largest launches of a frame; H = 5,7,8,10,13 one-sided taps)",
                    "strict": {"algorithmic_bytes": strict_bytes, "achieved": strict_bytes / (dom_ms * 1e-3) / 1e9, "frac": strict_bytes / (dom_ms * 1e-3) / 1e9 / peak,
                               "frac_stream_full": strict_bytes / (dom_ms_busy * 1e-3) / 1e9 / peak,
                               "how": "4 B written per pixel of the level (the read hits L2: the source level was just written) over the launch duration"},
                    "read_write": {"algorithmic_bytes": dom_bytes, "achieved": achieved, "frac": achieved / peak,
                                   "how": "SURVEY 8d's secondary model A_pyr2: 4 B read + 4 B written per pixel (+ the seed plane for level 3)"},
                    "fp32": {"ops_per_pixel": [4 * t + 2 for t in taps], "frac": fp32_frac, "peak_lane_ops_per_s": fp32_peak,
                             "how": "fp32 lane-instructions of the five launches over their summed duration, against SMs x 128 x max SM clock"},
                    "traffic": BLUR_DRAM_BYTES, "launch_ms": dom_ms, "launch_ms_stream_full": dom_ms_busy,
                    "per_level_ms": [float(x) for x in lv_alone], "per_level_ms_stream_full": [float(x) for x in lv], "per_level_bytes": lvl_bytes,
                    "timing": "CUDA events on the stream each launch goes to (psift_debug_time_levels), frames one at a time; *_stream_full: all frame slots busy"}}

    if rank == 0:
        line = {"metric": "images_per_sec_1080p", "value": world * n_frames / dt_dev, "unit": "images/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt_dev / args.steps,
                "ms_per_frame": 1e3 * dt_dev / n_frames, "wall_ms_per_step": 1e3 * wall_dev / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": workload_config(n_oct),
                "config_detail": {"in_flight": IN_FLIGHT,
                                  "timing": "CUDA events joining all of the context's streams around the K steps (max over ranks); wall clock of the same region in wall_ms_per_step; stage times from CUDA events on the frame's stream",
                                  "features_per_frame": nf / n_frames, "descriptors_per_frame": nd / n_frames},
                "e2e": {"value": world * n_frames / dt_cxx, "unit": "images/s", "h2d_bytes_per_step": W * H * BATCH,
                        "d2h_bytes_per_step": int(d2h_per_step),
                        "api": "the reference's C++ API: PopSift::enqueue(w, h, const unsigned char*) from pageable caller memory (private copy inside, as its contract says) and SiftJob::get() -> FeaturesHost, %d caller threads per GPU, %d jobs outstanding each; wall clock inside psift_cxx_stream, max over ranks" % (CXX_PRODUCERS, CXX_WINDOW),
                        "features_per_frame": cxx_feat / n_frames, "descriptors_per_frame": cxx_desc / n_frames,
                        "one_object": one_object,
                        "c_abi_pinned": {"value": world * n_frames / dt_e2e, "wall_value": world * n_frames / wall_e2e,
                                         "input": "page-locked host memory, psift_submit_pinned_u8 (uploaded straight from the caller's buffer), device-event time"},
                        "value_private_copy": world * n_frames / dt_copy,
                        "private_copy_note": "psift_submit_u8 = PopSift::enqueue()'s contract: pageable caller buffer, copied into the library's pinned staging before the upload (the caller may free it at once)",
                        "d2h_note": "result bytes actually queued for download (each frame copies a prefix sized from earlier frames, +25 %%, so that no host round trip sits between the kernels and the download); the frames hold %d bytes of results per step" % int((nf2 * (52 + 44) + nd2 * 512) / args.steps)},
                "gpu_launches": int(launches),
                "stage_ms": stage,
                "roofline": roofline,
                "cpu_baseline": cpu_baseline() if not args.no_cpu else {"value": None, "unit": "images/s", "cores": 1, "kind": "port", "sample": "skipped (--no-cpu)"},
                "tiled16k": tiled,
                "clocks": cs.summary()}
        print(json.dumps(line))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
