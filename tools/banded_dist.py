"""One giant image over the ranks of torch.distributed (one band per GPU), checked against the un-split run.
   torchrun --nproc-per-node N tools/banded_dist.py [W H]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch, torch.distributed as dist
from popsift_b200 import capi, synth, banded

rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
W, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (4096, 4096)
img = synth.gen(W, H, 5)
cfg = capi.default_config(downsampling=0)
with capi.Context(cfg, device=local) as ctx:
    mine, split, bounds = banded.extract_banded_dist(ctx, img, cfg)          # warm-up (allocations)
    torch.cuda.synchronize(); dist.barrier(); t0 = time.perf_counter()
    mine, split, bounds = banded.extract_banded_dist(ctx, img, cfg)
    torch.cuda.synchronize(); dist.barrier(); dt = time.perf_counter() - t0
    fs, ds = banded.digest(mine)
    counts = torch.tensor([mine.n_feat, mine.n_desc, fs, ds], device="cuda", dtype=torch.int64); dist.all_reduce(counts)
    if rank == 0:
        ctx.extract(img); t1 = time.perf_counter(); whole = ctx.extract(img); t2 = time.perf_counter()
        ok = (int(counts[0]), int(counts[1])) == (whole.n_feat, whole.n_desc)
        same = (int(counts[2]), int(counts[3])) == banded.digest(whole)
        print({"w": W, "h": H, "ranks": world, "split_octave": split, "bounds": bounds, "banded_s": round(dt, 4), "whole_1gpu_s": round(t2 - t1, 4),
               "features": int(counts[0]), "descriptors": int(counts[1]), "whole": (whole.n_feat, whole.n_desc), "counts_equal": ok, "checksums_equal": same})
dist.destroy_process_group()
