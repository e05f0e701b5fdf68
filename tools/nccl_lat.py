"""NCCL collective latencies at the C5 message sizes (torchrun, one rank per GPU): the numbers the engine's own
peer-memory exchange (tools/ipc_probe.cu) has to beat.  Device-timed, max over ranks."""
import json
import os

import torch
import torch.distributed as dist


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl")
    out = {"nranks": world}
    for name, n in (("allreduce_8B", 1), ("allreduce_48KB", 6000), ("allreduce_552KB", 69000), ("allreduce_144MB", 18_000_000)):
        x = torch.ones(n, dtype=torch.float64, device="cuda")
        for _ in range(5):
            dist.all_reduce(x)
        torch.cuda.synchronize(); dist.barrier()
        reps = 200 if n < 1_000_000 else 10
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            dist.all_reduce(x)
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) * 1e3 / reps], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out[name + "_us"] = round(float(t.item()), 2)
    n = 18_000_000 // world * world
    x = torch.ones(n, dtype=torch.float64, device="cuda")
    y = torch.empty(n // world, dtype=torch.float64, device="cuda")
    for _ in range(3):
        dist.reduce_scatter_tensor(y, x)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        dist.reduce_scatter_tensor(y, x)
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / 10], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    out["reduce_scatter_144MB_us"] = round(float(t.item()), 2)
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
