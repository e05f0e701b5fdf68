"""Per-chunk timeline of the pipelined batch call (B2BA_E2E_TRACE=1), on one GPU or on all ranks of a torchrun launch at
once (every rank uploads its own 1.18 GB per call: shows what the shared host side does to the H2D stream).
    B2BA_E2E_TRACE=1 python tools/e2e_trace.py
    B2BA_E2E_TRACE=1 python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/e2e_trace.py"""
import os
import sys
import time
from types import SimpleNamespace

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from visualodometry_b200 import BAConfig, BundleAdjuster, synth  # noqa: E402

rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
if rank not in (0, world - 1):
    os.environ["B2BA_E2E_TRACE"] = "0"          # two ranks print their timelines, all ranks run
b = synth.make_batch(4096, seed0=rank * 4096, distinct=64)
ba = BundleAdjuster(bench.K_of(b.intr), BAConfig(max_iters=10, device=local))
pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
hb = SimpleNamespace(poses=pin(b.poses), points=pin(b.points), obs_cam=pin(b.obs_cam), obs_pt=pin(b.obs_pt),
                     obs_uv=pin(b.obs_uv), fixed_cam=pin(b.fixed_cam), cam_off=b.cam_off, pt_off=b.pt_off, obs_off=b.obs_off)
out = (pin(np.empty_like(b.poses, dtype=np.float64)), pin(np.empty_like(b.points, dtype=np.float64)))
for rep in range(3):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t = time.perf_counter()
    ba.optimize_batch(hb, out=out)
    torch.cuda.synchronize()
    print("rank %d call %d: %.2f ms" % (rank, rep, 1e3 * (time.perf_counter() - t)), file=sys.stderr, flush=True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
