"""Rate of the C++ classes (PopSift::enqueue / SiftJob::get) at 1080p with one process per GPU, all running at once:
what the classes leg of `bench.py --gpus N` measures, without the other legs.  TEST INFRASTRUCTURE.
   python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/cxx_rate_ranks.py [frames]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
from popsift_b200 import capi, synth

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
dist.init_process_group("gloo")
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
imgs = [synth.gen(1920, 1080, 1000 + i) for i in range(16)]
capi.cxx_stream(imgs, 50, 200, first_device=local, n_devices=1, n_producers=2, window=6)      # contexts, pools, clocks
dist.barrier()
t0 = time.perf_counter()
r = capi.cxx_stream(imgs, 100, frames, first_device=local, n_devices=1, n_producers=2, window=6)
t = torch.tensor([r["seconds"]], dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print("ranks %d  PSIFT_BLOCKING_SYNC=%s  classes: %.1f images/s (slowest rank %.3f s for %d frames each)"
          % (world, os.environ.get("PSIFT_BLOCKING_SYNC", "0"), world * frames / t.item(), t.item(), frames), flush=True)
dist.destroy_process_group()
