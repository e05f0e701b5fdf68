"""C5 strong scaling on the GPUs of one box (torchrun): both visibilities, ms per LM iteration and its phase split.
python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/c5_scaling.py  -> one JSON line."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    out = {"n_gpus": world}
    for mode in (sys.argv[1:] or ["local", "random"]):
        out[mode] = bench.c5_measure(mode, "auto", 20, 3, 2, rank, world, local)
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
