"""Run under torchrun on >= 2 GPUs: the multi-rank large path (S sharded by block row, replicated small state, exchanges
through peer memory: csrc/ba_large.cuh) must give the single-GPU result for every solver (banded Cholesky, PCG, dense
DMMA Cholesky).  Rank 0 prints one line per case and exits non-zero
on a mismatch.   python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/multi_gpu_check.py"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from visualodometry_b200 import BAConfig, BundleAdjuster, synth  # noqa: E402


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank = dist.get_rank()
    bad = 0
    for mode, nc, npts, solver in (("local", 120, 12000, "auto"), ("random", 120, 12000, "pcg"), ("local", 400, 60000, "auto"),
                                   ("random", 300, 40000, "pcg"), ("random", 120, 12000, "dense")):
        p = synth.make_large(7, n_cams=nc, n_pts=npts, mode=mode)
        K = np.array([[p.intr[0], 0, p.intr[2]], [0, p.intr[1], p.intr[3]], [0, 0, 1.0]])
        cfg = BAConfig(max_iters=6, solver=solver, device=local)
        single = BundleAdjuster(K, cfg)
        single.set_problem(p.poses, p.points, p.obs_cam, p.obs_pt, p.obs_uv, p.fixed_cam)
        s1 = single.solve()[0]
        po1, x1 = single.get()
        single.close()
        multi = BundleAdjuster(K, cfg).init_distributed()
        multi.set_problem(p.poses, p.points, p.obs_cam, p.obs_pt, p.obs_uv, p.fixed_cam)
        s2 = multi.solve()[0]
        po2, x2 = multi.get()
        multi.close()
        dp = np.abs(po1 - po2).max()
        dx = (np.linalg.norm(x1 - x2, axis=1) / np.maximum(1, np.linalg.norm(x1, axis=1))).max()
        dc = abs(s1.final_cost - s2.final_cost) / s1.final_cost
        ok = dp < 1e-7 and dx < 1e-7 and dc < 1e-9 and (s1.accepted, s1.rejected) == (s2.accepted, s2.rejected)
        if rank == 0:
            print(f"{mode} nc={nc} np={npts} solver={solver}: ranks={dist.get_world_size()} |dpose|={dp:.2e} |dX|rel={dx:.2e} dcost={dc:.2e} "
                  f"accepted {s1.accepted}/{s2.accepted} pcg {s1.pcg_iters}/{s2.pcg_iters} -> {'OK' if ok else 'MISMATCH'}", flush=True)
        bad += 0 if ok else 1
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
